/* Link-time stubs for the oracle build of the UNMODIFIED reference (oracle/Makefile).
 * ARPACK and UMFPACK are absent from this image; the reference's eigen-solvers
 * (src/linalg.c:328-1000) reference these symbols.  None of them is on the apply path
 * (SURVEY.md §8c), so reaching one is a bug in a test driver: abort loudly.
 * Test infrastructure, not product. */
#include <stdio.h>
#include <stdlib.h>

#define ORACLE_STUB(name)                                                   \
  void name(void) {                                                         \
    fprintf(stderr, "oracle: %s is not available in this build\n", #name);  \
    abort();                                                                \
  }

ORACLE_STUB(dnaupd_c)
ORACLE_STUB(dneupd_c)
ORACLE_STUB(bfLuCsrRealNew)
ORACLE_STUB(bfLuCsrRealInit)
ORACLE_STUB(bfLuCsrRealSolveVec)
ORACLE_STUB(bfLuCsrRealDeinitAndDealloc)
