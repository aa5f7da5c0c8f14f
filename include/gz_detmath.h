/*
 * gz_detmath.h — a reproducible libm subset: sin, cos, atan2 in IEEE-754 double.
 *
 * Why: the step's only transcendental calls are atan2 (hinge angle,
 * deps/opende/src/joints/joint.cpp:391-432, called from hinge.cpp:65-77 and
 * dJointGetHingeAngle) and sin/cos (plugin-side gait generators).  glibc's and CUDA's
 * libm differ in the last bit on some arguments, and an actuated robot bouncing on the
 * ground amplifies one ulp to millimetres within a second of simulated time.  Every
 * function here uses only +, -, *, /, comparisons and table constants, in a fixed order,
 * so it returns the SAME BITS on the host (compiled with -ffp-contract=off) and on the
 * device (compiled with -fmad=false).  libgzcuda uses it on the device; the parity
 * oracle links the reference's opende against it instead of glibc's libm for these three
 * symbols (oracle/shim/detmath_shim.h).  Accuracy: atan2 < 1.5 ulp, sin/cos < 1 ulp over the tested ranges
 * (tests/test_detmath.py); argument range of sin/cos: |x| < 1e5.
 *
 * Constants come from tools/gen_detmath_tables.py (mpmath, 60 digits).
 */
#ifndef GZ_DETMATH_H
#define GZ_DETMATH_H

#if defined(__CUDACC__)
#define GZDM_HD __host__ __device__ static inline
#else
#define GZDM_HD static inline
#endif

/* atan(k/8) = hi + lo */
GZDM_HD void gzdm_atan_tab(int k, double *hi, double *lo) {
  switch (k) {
    case 0: *hi = 0.0; *lo = 0.0; break;
    case 1: *hi = 0x1.fd5ba9aac2f6ep-4; *lo = -0x1.cd37686760c17p-59; break;
    case 2: *hi = 0x1.f5b75f92c80ddp-3; *lo = 0x1.8ab6e3cf7afbdp-57; break;
    case 3: *hi = 0x1.6f61941e4def1p-2; *lo = -0x1.c63aae6f6e918p-56; break;
    case 4: *hi = 0x1.dac670561bb4fp-2; *lo = 0x1.a2b7f222f65e2p-56; break;
    case 5: *hi = 0x1.1e00babdefeb4p-1; *lo = -0x1.928df287a668fp-58; break;
    case 6: *hi = 0x1.4978fa3269ee1p-1; *lo = 0x1.2419a87f2a458p-56; break;
    case 7: *hi = 0x1.700a7c5784634p-1; *lo = -0x1.8c34d25aadef6p-56; break;
    default: *hi = 0x1.921fb54442d18p-1; *lo = 0x1.1a62633145c07p-55; break;
  }
}

/* atan(t) for 0 <= t <= 1, returned as hi + lo (hi carries the value, lo the tail) */
GZDM_HD void gzdm_atan01(double t, double *rhi, double *rlo) {
  int k = (int)(t * 8.0 + 0.5);
  double c = (double)k * 0.125;
  double u = (k == 0) ? t : (t - c) / (1.0 + t * c);   /* |u| <= 1/16 */
  double z = u * u;
  /* atan(u) = u - u^3/3 + u^5/5 - ...  (terms to u^17; truncation < 2^-80) */
  double p = -1.0 / 15.0 + z * (1.0 / 17.0);
  p = 1.0 / 13.0 + z * p;
  p = -1.0 / 11.0 + z * p;
  p = 1.0 / 9.0 + z * p;
  p = -1.0 / 7.0 + z * p;
  p = 1.0 / 5.0 + z * p;
  p = -1.0 / 3.0 + z * p;
  double hi, lo;
  gzdm_atan_tab(k, &hi, &lo);
  *rhi = hi;
  *rlo = u + (lo + u * (z * p));
}

GZDM_HD double gzdet_atan2(double y, double x) {
  const double pi_hi = 0x1.921fb54442d18p+1, pi_lo = 0x1.1a62633145c07p-53;
  const double pio2_hi = 0x1.921fb54442d18p+0, pio2_lo = 0x1.1a62633145c07p-54;
  if (x != x || y != y) return x + y;
  double ax = x < 0 ? -x : x, ay = y < 0 ? -y : y;
  int xneg = (x < 0) || (x == 0 && 1.0 / x < 0);
  int yneg = (y < 0) || (y == 0 && 1.0 / y < 0);
  double r;
  if (ay == 0) {
    r = xneg ? pi_hi : 0.0;
  } else if (ax == 0) {
    r = pio2_hi;
  } else if (ax > 1.7e308 || ay > 1.7e308) {          /* infinities */
    if (ax > 1.7e308 && ay > 1.7e308) r = xneg ? 3.0 * (pio2_hi * 0.5) : pio2_hi * 0.5;
    else if (ax > 1.7e308) r = xneg ? pi_hi : 0.0;
    else r = pio2_hi;
  } else {
    int swap = ay > ax;
    double t = swap ? ax / ay : ay / ax;
    double hi, lo;
    gzdm_atan01(t, &hi, &lo);
    if (swap) {            /* pi/2 - a */
      lo = pio2_lo - lo;
      hi = pio2_hi - hi;
    }
    if (xneg) {            /* pi - a */
      lo = pi_lo - lo;
      hi = pi_hi - hi;
    }
    r = hi + lo;
  }
  return yneg ? -r : r;
}

/* kernels on |r| <= pi/4 (+ tail rl) */
GZDM_HD double gzdm_ksin(double r, double rl) {
  double z = r * r;
  double p = -1.0 / 1307674368000.0 + z * (1.0 / 355687428096000.0);   /* -1/15! , 1/17! */
  p = 1.0 / 6227020800.0 + z * p;      /* 1/13! */
  p = -1.0 / 39916800.0 + z * p;       /* -1/11! */
  p = 1.0 / 362880.0 + z * p;          /* 1/9! */
  p = -1.0 / 5040.0 + z * p;           /* -1/7! */
  p = 1.0 / 120.0 + z * p;             /* 1/5! */
  p = -1.0 / 6.0 + z * p;              /* -1/3! */
  return r + (rl + r * (z * p));
}
GZDM_HD double gzdm_kcos(double r, double rl) {
  double z = r * r;
  double p = -1.0 / 6402373705728000.0 + z * (1.0 / 2432902008176640000.0);   /* -1/18!, 1/20! */
  p = 1.0 / 20922789888000.0 + z * p;  /* 1/16! */
  p = -1.0 / 87178291200.0 + z * p;    /* -1/14! */
  p = 1.0 / 479001600.0 + z * p;       /* 1/12! */
  p = -1.0 / 3628800.0 + z * p;        /* -1/10! */
  p = 1.0 / 40320.0 + z * p;           /* 1/8! */
  p = -1.0 / 720.0 + z * p;            /* -1/6! */
  p = 1.0 / 24.0 + z * p;              /* 1/4! */
  double hz = 0.5 * z;
  double w = 1.0 - hz;
  return w + (((1.0 - w) - hz) + (z * (z * p) - r * rl));
}

/* r + rl = x - n*pi/2, n returned */
GZDM_HD int gzdm_reduce(double x, double *r, double *rl) {
  const double two_over_pi = 0x1.45f306dc9c883p-1;
  const double P1 = 0x1.921fb54400000p+0, P2 = 0x1.0b4611a600000p-34, P3 = 0x1.3198a2e037073p-69;
  double fn = x * two_over_pi;
  fn = (fn >= 0) ? (double)(long long)(fn + 0.5) : -(double)(long long)(-fn + 0.5);
  double a = x - fn * P1;          /* exact: fn*P1 has <= 53 bits for |fn| < 2^20 */
  double b = fn * P2;              /* exact */
  double hi = a - b;
  double lo = ((a - hi) - b) - fn * P3;
  *r = hi + lo;
  *rl = (hi - *r) + lo;
  return (int)((long long)fn & 3);
}

GZDM_HD double gzdet_sin(double x) {
  if (x != x) return x;
  double r, rl;
  int n = gzdm_reduce(x, &r, &rl);
  switch (n) {
    case 0: return gzdm_ksin(r, rl);
    case 1: return gzdm_kcos(r, rl);
    case 2: return -gzdm_ksin(r, rl);
    default: return -gzdm_kcos(r, rl);
  }
}

GZDM_HD double gzdet_cos(double x) {
  if (x != x) return x;
  double r, rl;
  int n = gzdm_reduce(x, &r, &rl);
  switch (n) {
    case 0: return gzdm_kcos(r, rl);
    case 1: return -gzdm_ksin(r, rl);
    case 2: return -gzdm_kcos(r, rl);
    default: return gzdm_ksin(r, rl);
  }
}

#endif /* GZ_DETMATH_H */
