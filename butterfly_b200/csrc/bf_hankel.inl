// bf_hankel.inl — H0^(1) and H1^(1) in fp64, written once for the device (bf_helm2.cu) and for a host
// build that tests/test_hankel_host.py checks against scipy without a GPU.  The includer defines
//   BF_HANKEL_FN     function qualifiers      (__device__   |  nothing)
//   BF_HANKEL_TABLE  table storage qualifiers (__constant__ |  static const)
// and provides double2 / make_double2 / sincos.
// ---- Hankel functions ---------------------------------------------------------------------------------
// CUDA's double-precision j0 / y0 / j1 / y1 are only good to an ABSOLUTE error of ~5e-12 for |x| >= 8
// (CUDA math API accuracy tables) — three digits short of the 1e-12 the apply is held to, since |H0(kr)| is
// ~0.02 at the distances that dominate a butterfly.  H0 therefore follows the reference's own
// Chebyshev evaluation (src/bessel.c:14-120,197-262: the GSL / SLATEC series bj0, by0 on [0, 4] and the
// amplitude-phase series bm0, bth0 beyond, with the phase added through the angle-sum identities so that
// no pi/4 is subtracted from a large argument).  H1 uses the Cephes rational approximations (j1.c: RP/RQ,
// YP/YQ on [0, 5], PP/PQ, QP/QQ in 25/x^2 beyond), the same phase treatment.  Both agree with glibc's
// j0 / y0 / j1 / y1 — what the reference calls in a _DEFAULT_SOURCE build, src/bessel.c:290-308 — to 1e-15.
BF_HANKEL_TABLE double cBj0[13] = {0.100254161968939137, -0.665223007764405132, 0.248983703498281314, -0.0332527231700357697,
                                0.0023114179304694015, -0.0000991127741995080, 0.0000028916708643998, -0.0000000612108586630,
                                0.0000000009838650793, -0.0000000000124235515, 0.0000000000001265433, -0.0000000000000010619,
                                0.0000000000000000074};
BF_HANKEL_TABLE double cBy0[13] = {-0.011277839392865573, -0.128345237560420350, -0.104378847997942490, 0.023662749183969695,
                                -0.002090391647700486, 0.000103975453939057, -0.000003369747162423, 0.000000077293842676,
                                -0.000000001324976772, 0.000000000017648232, -0.000000000000188105, 0.000000000000001641,
                                -0.000000000000000011};
BF_HANKEL_TABLE double cBm0[21] = {0.09284961637381644, -0.00142987707403484, 0.00002830579271257, -0.00000143300611424,
                                0.00000012028628046, -0.00000001397113013, 0.00000000204076188, -0.00000000035399669,
                                0.00000000007024759, -0.00000000001554107, 0.00000000000376226, -0.00000000000098282,
                                0.00000000000027408, -0.00000000000008091, 0.00000000000002511, -0.00000000000000814,
                                0.00000000000000275, -0.00000000000000096, 0.00000000000000034, -0.00000000000000012,
                                0.00000000000000004};
BF_HANKEL_TABLE double cBth0[24] = {-0.24639163774300119, 0.001737098307508963, -0.000062183633402968, 0.000004368050165742,
                                 -0.000000456093019869, 0.000000062197400101, -0.000000010300442889, 0.000000001979526776,
                                 -0.000000000428198396, 0.000000000102035840, -0.000000000026363898, 0.000000000007297935,
                                 -0.000000000002144188, 0.000000000000663693, -0.000000000000215126, 0.000000000000072659,
                                 -0.000000000000025465, 0.000000000000009229, -0.000000000000003448, 0.000000000000001325,
                                 -0.000000000000000522, 0.000000000000000210, -0.000000000000000087, 0.000000000000000036};

// sum' c_j T_j(x) by Clenshaw's recurrence, the c_0 term halved (chebEvalGsl, src/bessel.c:122-135)
template <int N> BF_HANKEL_FN inline double chebSeries(const double (&c)[N], double x) {
  double d = 0.0, dd = 0.0;
  const double x2 = 2.0 * x;
#pragma unroll
  for (int j = N - 1; j >= 1; --j) {
    const double t = d;
    d = fma(x2, d, c[j] - dd);
    dd = t;
  }
  return fma(x, d, 0.5 * c[0] - dd);
}

// (cos, sin)(x + phase0 + eps) for phase0 = -pi/4 (quarter = 1) or -3 pi/4 (quarter = 3), |eps| small
BF_HANKEL_FN inline void shiftedSinCos(double x, double eps, int quarter, double *cp, double *sp) {
  double sx, cx, se, ce;
  sincos(x, &sx, &cx);
  sincos(eps, &se, &ce);
  const double r = 0.70710678118654752440;
  // cos(x - pi/4) = (cx + sx) r, sin(x - pi/4) = (sx - cx) r; cos(x - 3pi/4) = (sx - cx) r, sin(x - 3pi/4) = -(sx + cx) r
  const double c0 = quarter == 1 ? (cx + sx) * r : (sx - cx) * r;
  const double s0 = quarter == 1 ? (sx - cx) * r : -(sx + cx) * r;
  *cp = c0 * ce - s0 * se;
  *sp = s0 * ce + c0 * se;
}

// H0^(1)(x) = J0 + i Y0, x > 0
BF_HANKEL_FN inline double2 hankel0(double x) {
  if (x <= 4.0) {
    const double z = 0.125 * x * x - 1.0;
    const double j = x < 2.9802322387695312e-08 ? 1.0 : chebSeries(cBj0, z); // 2 sqrt(eps)
    const double y = 0.63661977236758134308 * (log(x) - 0.69314718055994530942) * j + 0.375 + chebSeries(cBy0, z);
    return make_double2(j, y);
  }
  const double z = 32.0 / (x * x) - 1.0;
  const double amp = (0.75 + chebSeries(cBm0, z)) / sqrt(x);
  double cp, sp;
  shiftedSinCos(x, chebSeries(cBth0, z) / x, 1, &cp, &sp);
  return make_double2(amp * cp, amp * sp);
}

BF_HANKEL_FN inline double polyN(const double *c, int n, double z, double lead) {
  double p = lead; // lead = 1 for the monic denominators (Cephes p1evl), 0 otherwise
  for (int i = 0; i < n; ++i) p = fma(p, z, c[i]);
  return p;
}
BF_HANKEL_TABLE double cPP[7] = {7.62125616208173112003e-04, 7.31397056940917570436e-02, 1.12719608129684925192e+00,
                              5.11207951146807644818e+00, 8.42404590141772420927e+00, 5.21451598682361504063e+00,
                              1.00000000000000000254e+00};
BF_HANKEL_TABLE double cPQ[7] = {5.71323128072548699714e-04, 6.88455908754495404082e-02, 1.10514232634061696926e+00,
                              5.07386386128601488557e+00, 8.39985554327604159757e+00, 5.20982848682361821619e+00,
                              9.99999999999999997461e-01};
BF_HANKEL_TABLE double cQP[8] = {5.10862594750176621635e-02, 4.98213872951233449420e+00, 7.58238284132545283818e+01,
                              3.66779609360150777800e+02, 7.10856304998926107277e+02, 5.97489612400613639965e+02,
                              2.11688757100572135698e+02, 2.52070205858023719784e+01};
BF_HANKEL_TABLE double cQQ[7] = {7.42373277035675149943e+01, 1.05644886038262816351e+03, 4.98641058337653607651e+03,
                              9.56231892404756170795e+03, 7.99704160447350683650e+03, 2.82619278517639096600e+03,
                              3.36093607810698293419e+02};
BF_HANKEL_TABLE double cRP[4] = {-8.99971225705559398224e+08, 4.52228297998194034323e+11, -7.27494245221818276015e+13,
                              3.68295732863852883286e+15};
BF_HANKEL_TABLE double cRQ[8] = {6.20836478118054335476e+02, 2.56987256757748830383e+05, 8.35146791431949253037e+07,
                              2.21511595479792499675e+10, 4.74914122079991414898e+12, 7.84369607876235854894e+14,
                              8.95222336184627338078e+16, 5.32278620332680085395e+18};
BF_HANKEL_TABLE double cYP[6] = {1.26320474790178026440e+09, -6.47355876379160291031e+11, 1.14509511541823727583e+14,
                              -8.12770255501325109621e+15, 2.02439475713594898196e+17, -7.78877196265950026825e+17};
BF_HANKEL_TABLE double cYQ[8] = {5.94301592346128195359e+02, 2.35564092943068577943e+05, 7.34811944459721705660e+07,
                              1.87601316108706159478e+10, 3.88231277496238566008e+12, 6.20557727146953693363e+14,
                              6.87141087355300489866e+16, 3.97270608116560655612e+18};

// H1^(1)(x) = J1 + i Y1, x > 0
BF_HANKEL_FN inline double2 hankel1(double x) {
  if (x <= 5.0) {
    const double z = x * x;
    const double j = polyN(cRP, 4, z, 0.0) / polyN(cRQ, 8, z, 1.0) * x * (z - 1.46819706421238932572e+01) * (z - 4.92184563216946036703e+01);
    const double y = x * (polyN(cYP, 6, z, 0.0) / polyN(cYQ, 8, z, 1.0)) + 0.63661977236758134308 * (j * log(x) - 1.0 / x);
    return make_double2(j, y);
  }
  const double w = 5.0 / x, z = w * w;
  const double p = polyN(cPP, 7, z, 0.0) / polyN(cPQ, 7, z, 0.0);
  const double q = polyN(cQP, 8, z, 0.0) / polyN(cQQ, 7, z, 1.0);
  double cp, sp;
  shiftedSinCos(x, 0.0, 3, &cp, &sp);
  const double amp = 0.79788456080286535588 / sqrt(x); // sqrt(2 / pi)
  return make_double2((p * cp - w * q * sp) * amp, (p * sp + w * q * cp) * amp);
}

