"""numpy interpreter of the compiled plan tables (include/bf_cuda_debug.h).  TEST CODE: it walks
stages / items / tiles exactly as the kernels do, on NaN-initialised buffers, so that any row a
kernel would read before it is written, any aliasing of live workspace rows and any mistake in
tile addressing or pool packing shows up without a GPU."""
import ctypes as C

import numpy as np

from butterfly_b200 import capi


class Tables:
    def __init__(self, raw, op=capi.OP_N, num_queues=148 * 8, shard=None):
        self.lib = capi.load_library()
        self.raw = raw
        self._desc = raw.desc(shard=shard)
        t = C.POINTER(capi.DebugTables)()
        capi.check(self.lib.bfCudaDebugCompile(C.byref(self._desc), op, num_queues, C.byref(t)), "bfCudaDebugCompile")
        self._t = t
        d = t.contents
        self.d = d
        nS, Q = d.numStages, d.numQueues
        arr = lambda p, n, dt: np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_ubyte)), shape=(n * np.dtype(dt).itemsize,)).view(dt).copy() if n else np.zeros(0, dt)
        self.stage_item = arr(d.stageItemBegin, nS + 1, np.int64)
        self.stage_tile = arr(d.stageTileBegin, nS + 1, np.int64)
        self.items = arr(d.items, d.numItems, capi.ITEM_DTYPE)
        self.tiles = arr(d.tiles, d.numTiles, capi.TILE_DTYPE)
        self.tile_pool16 = arr(d.tilePool16, d.numTiles, np.uint64)
        self.q_item = arr(d.qItemBegin, nS * (Q + 1), np.uint32).reshape(nS, Q + 1)
        self.q_tile = arr(d.qTileBegin, nS * (Q + 1), np.uint32).reshape(nS, Q + 1)
        self.q_pool16 = arr(d.qPool16, nS * (Q + 1), np.uint64).reshape(nS, Q + 1)
        self.tile_op = arr(d.tileOp, d.numTiles, np.int64)
        self.pool = arr(d.pool, int(d.pool16) * 2, np.float64) if d.pool else np.zeros(0)
        self.num_stages, self.num_queues = nS, Q
        self.rows_in, self.rows_out, self.ws_rows = d.rowsIn, d.rowsOut, d.wsRows
        self.cplx = d.dtype == capi.COMPLEX_F64
        self.ws_arena = (int(d.wsArena0Rows), int(d.wsArena1Rows))
        self.item_peers = arr(d.itemPeers, d.numItems, np.uint32) if d.itemPeers else None
        self.recv_mask = arr(d.stageRecvMask, nS, np.uint32) if d.stageRecvMask else None
        self.send_mask = arr(d.stageSendMask, nS, np.uint32) if d.stageSendMask else None
        self.fused = bool(d.fusedEnabled)
        if self.fused:
            self.f_items = arr(d.fusedItems, d.fusedNumItems, capi.ITEM_DTYPE)
            self.f_item_deps = arr(d.fusedItemDeps, d.fusedNumItems, capi.ITEM_DEP_DTYPE)
            self.f_tiles = arr(d.fusedTiles, d.fusedNumTiles, capi.TILE_DTYPE)
            self.f_tile_deps = arr(d.fusedTileDeps, d.fusedNumTiles, capi.TILE_DEP_DTYPE)
            self.f_q_item = arr(d.fusedQItemBegin, Q + 1, np.uint32)
            self.f_q_tile = arr(d.fusedQTileBegin, Q + 1, np.uint32)
            self.f_ws_rows, self.f_num_counters = int(d.fusedWsRows), int(d.fusedNumCounters)

    def close(self):
        if self._t:
            self.lib.bfCudaDebugFree(self._t)
            self._t = None

    def new_space(self, x):
        """[X, Y, workspace] with everything but X NaN-initialised."""
        dt = np.complex128 if self.cplx else np.float64
        x = np.asarray(x, dtype=dt)
        if x.ndim == 1:
            x = x[:, None]
        nrhs = x.shape[1]
        nan = np.nan * (1 + 1j) if self.cplx else np.nan
        return [x, np.full((self.rows_out, nrhs), nan, dt), np.full((self.ws_rows + 2, nrhs), nan, dt)]

    def run_stage(self, space, s):
        dt = space[1].dtype
        nrhs = space[0].shape[1]
        e = 2 if self.cplx else 1
        i0, i1 = self.stage_item[s], self.stage_item[s + 1]
        t0 = self.stage_tile[s]
        # an item's reads see the state before this stage (no intra-stage dependencies allowed)
        results = []
        for it in self.items[i0:i1]:
            rows = int(it["rows"])
            acc = np.zeros((rows, nrhs), dt)
            for t in range(t0 + int(it["tileBegin"]), t0 + int(it["tileBegin"]) + int(it["numTiles"])):
                tl = self.tiles[t]
                assert int(tl["rows"]) == rows
                sp, r = int(tl["inAddr"]) >> 30, int(tl["inAddr"]) & capi.ROW_MASK
                kc = int(tl["kc"])
                p0 = int(self.tile_pool16[t]) * 2
                if kc == capi.KC_IDENTITY:
                    acc += space[sp][r:r + rows]
                elif kc == capi.KC_DIAG:
                    acc += self.pool[p0:p0 + rows, None] * space[sp][r:r + rows]
                else:
                    a = self.pool[p0:p0 + rows * kc * e]
                    a = a.view(np.complex128) if self.cplx else a
                    acc += a.reshape(rows, kc) @ space[sp][r:r + kc]
            results.append(acc)
        for it, acc in zip(self.items[i0:i1], results):
            sp, r = int(it["outAddr"]) >> 30, int(it["outAddr"]) & capi.ROW_MASK
            assert sp != capi.SPACE_X
            rows = int(it["rows"])
            if int(it["flags"]) & capi.ITEM_ACCUMULATE:
                space[sp][r:r + rows] += acc
            else:
                space[sp][r:r + rows] = acc

    def apply(self, x):
        space = self.new_space(x)
        for s in range(self.num_stages):
            self.run_stage(space, s)
        return space[1]

    def apply_fused(self, x, order="reverse", seed=0):
        """The fused single-launch tables (bf_plan.hpp FusedTables) run the way bfGemvStage<., true> runs them:
        every queue walks its own item list in order, a tile may only be consumed once its operand buffer's
        counter has reached its total, an accumulating item may only store once its counter has reached the
        earlier stages' count, and a store bumps the item's counter.  The SCHEDULE is adversarial on purpose
        (queues visited in reverse / random order, each running until it blocks): a missing dependency shows
        up as a NaN read from the NaN-initialised flat workspace or as a wrong sum, a cyclic one as a deadlock.
        Returns (Y, number of times a queue had to wait)."""
        assert self.fused
        dt = np.complex128 if self.cplx else np.float64
        x = np.asarray(x, dtype=dt)
        if x.ndim == 1:
            x = x[:, None]
        nrhs = x.shape[1]
        nan = np.nan * (1 + 1j) if self.cplx else np.nan
        space = [x, np.full((self.rows_out, nrhs), nan, dt), np.full((self.f_ws_rows + 2, nrhs), nan, dt)]
        cnt = np.zeros(max(self.f_num_counters, 1), np.int64)
        e = 2 if self.cplx else 1
        Q = self.num_queues
        cur = self.f_q_item[:-1].astype(np.int64).copy()
        end = self.f_q_item[1:].astype(np.int64)
        cpool = np.zeros(Q, np.int64)  # consumer pool cursor per queue, in doubles
        rng = np.random.default_rng(seed)
        waits = 0
        NO = capi.NO_COUNTER

        def try_item(q):
            """Run the queue's next item if all of its waits are satisfied; False if it has to wait."""
            i = cur[q]
            it, dep = self.f_items[i], self.f_item_deps[i]
            tb, nt, rows = int(it["tileBegin"]), int(it["numTiles"]), int(it["rows"])
            for t in range(tb, tb + nt):
                td = self.f_tile_deps[t]
                if int(td["idx"]) != NO and cnt[int(td["idx"])] < int(td["total"]):
                    return False
            acc_flag = int(it["flags"]) & capi.ITEM_ACCUMULATE
            if acc_flag and int(dep["sig"]) != NO and cnt[int(dep["sig"])] < int(dep["waitTarget"]):
                return False
            acc = np.zeros((rows, nrhs), dt)
            p = cpool[q]
            for t in range(tb, tb + nt):
                tl = self.f_tiles[t]
                kc = int(tl["kc"])
                if kc == capi.KC_JUMP:
                    p = ((int(tl["rows"]) << 32) | int(tl["inAddr"])) * 2
                    continue
                assert int(tl["rows"]) == rows
                sp, r = int(tl["inAddr"]) >> 30, int(tl["inAddr"]) & capi.ROW_MASK
                if kc == capi.KC_IDENTITY:
                    acc += space[sp][r:r + rows]
                elif kc == capi.KC_DIAG:
                    acc += self.pool[p:p + rows, None] * space[sp][r:r + rows]
                    p += ((rows * 8 + 15) // 16) * 2
                else:
                    a = self.pool[p:p + rows * kc * e]
                    a = a.view(np.complex128) if self.cplx else a
                    acc += a.reshape(rows, kc) @ space[sp][r:r + kc]
                    p += ((rows * kc * 8 * e + 15) // 16) * 2
            cpool[q] = p
            if not (int(it["flags"]) & capi.ITEM_JUMP):
                sp, r = int(it["outAddr"]) >> 30, int(it["outAddr"]) & capi.ROW_MASK
                assert sp != capi.SPACE_X
                if acc_flag:
                    space[sp][r:r + rows] += acc
                else:
                    space[sp][r:r + rows] = acc
                if int(dep["sig"]) != NO:
                    cnt[int(dep["sig"])] += 1
            cur[q] += 1
            return True

        while True:
            live = [q for q in range(Q) if cur[q] < end[q]]
            if not live:
                break
            if order == "reverse":
                live.reverse()
            elif order == "random":
                rng.shuffle(live)
            progressed = False
            for q in live:
                burst = int(rng.integers(1, 6)) if order == "random" else 1 << 30
                while cur[q] < end[q] and burst > 0:
                    if not try_item(q):
                        waits += 1
                        break
                    progressed = True
                    burst -= 1
            assert progressed, "deadlock: every live queue waits on a counter"
        # every counter ends at the total its readers wait for
        for td in self.f_tile_deps:
            if int(td["idx"]) != NO:
                assert cnt[int(td["idx"])] == int(td["total"])
        return space[1], waits

    def check_queues(self):
        """Queues partition every stage's items in order; each queue's pool offset is where its
        first tile lives; tiles are packed back to back in consumption order."""
        for s in range(self.num_stages):
            nI = self.stage_item[s + 1] - self.stage_item[s]
            nT = self.stage_tile[s + 1] - self.stage_tile[s]
            qi, qt, qp = self.q_item[s], self.q_tile[s], self.q_pool16[s]
            assert qi[0] == 0 and qi[-1] == nI and np.all(np.diff(qi.astype(np.int64)) >= 0)
            assert qt[0] == 0 and qt[-1] == nT and np.all(np.diff(qt.astype(np.int64)) >= 0)
            items = self.items[self.stage_item[s]:self.stage_item[s + 1]]
            tp = self.tile_pool16[self.stage_tile[s]:self.stage_tile[s + 1]]
            for q in range(self.num_queues):
                if qi[q] < nI:
                    assert qt[q] == items[qi[q]]["tileBegin"]
                if qt[q] < nT:
                    assert qp[q] == tp[qt[q]]
            # back-to-back packing
            tl = self.tiles[self.stage_tile[s]:self.stage_tile[s + 1]]
            es = 16 if self.cplx else 8
            kc = tl["kc"].astype(np.int64)
            rows = tl["rows"].astype(np.int64)
            b16 = np.where(kc == capi.KC_IDENTITY, 0, np.where(kc == capi.KC_DIAG, (rows * 8 + 15) // 16, (rows * kc * es + 15) // 16))
            if nT:
                assert np.all(np.diff(tp.astype(np.int64)) == b16[:-1])
