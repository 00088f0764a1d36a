// Epsilon-greedy and Thompson-sampling routers, batched: T interleaved route/send_feedback calls per launch.
//
// Replaces components/routers/epsilon-greedy/EpsilonGreedy.py:35-136 and
// components/routers/thompson-sampling/ThompsonSampling.py:35-115 of the reference.  Decisions must be
// bit-identical to the reference under a seed, so both MT19937 streams it draws from are reproduced word for word
// on the device:
//   - CPython `random` : init_by_array seeding, 53-bit doubles from two words, _randbelow by rejection
//                        (route: random.random(), random.choice(others); initial best_branch)
//   - NumPy legacy     : init_genrand seeding, masked-rejection bounded integers (tie break, np.random.choice),
//                        legacy beta = Johnk (a,b <= 1) or two Marsaglia-Tsang gammas fed by the cached polar
//                        Gaussian (np.random.beta)
// The bandit is sequential in time (route_t depends on feedback_{t-1}); this is a latency kernel: warp 0 runs it, all
// lanes executing the same instruction stream (the generator state lives in shared memory and is advanced
// uniformly), lanes = arms for the tie-break scan.  It has no roofline; bench.py reports calls/s.
#include <string.h>

#include "common.cuh"

namespace {

constexpr int MT_N = 624, MT_M = 397;
constexpr int kMaxArms = 1024;

struct MtState {
  uint32_t mt[MT_N];
  int pos;
  int has_gauss;
  double gauss;
};

struct RouterState {
  MtState py, np;
  long long succ[kMaxArms];
  long long tries[kMaxArms];
  int best;
  long long guard_hits;  // accept/reject comparisons that fell inside the libm guard band (resolved with cr_log)
  long long fallbacks;   // beta draws taken draw by draw because they outran the look-ahead window
  double last_values[kMaxArms];  // Thompson: the beta samples of the last route (branch_values of the reference class)
};

__host__ __device__ inline void mt_init_genrand(MtState* g, uint32_t s) {
  g->mt[0] = s;
  for (int i = 1; i < MT_N; ++i) g->mt[i] = 1812433253u * (g->mt[i - 1] ^ (g->mt[i - 1] >> 30)) + (uint32_t)i;
  g->pos = MT_N;
  g->has_gauss = 0;
  g->gauss = 0.0;
}

inline void mt_init_by_array(MtState* g, const uint32_t* key, int len) {
  mt_init_genrand(g, 19650218u);
  uint32_t* mt = g->mt;
  int i = 1, j = 0;
  for (int k = MT_N > len ? MT_N : len; k; --k) {
    mt[i] = (mt[i] ^ ((mt[i - 1] ^ (mt[i - 1] >> 30)) * 1664525u)) + key[j] + (uint32_t)j;
    ++i; ++j;
    if (i >= MT_N) { mt[0] = mt[MT_N - 1]; i = 1; }
    if (j >= len) j = 0;
  }
  for (int k = MT_N - 1; k; --k) {
    mt[i] = (mt[i] ^ ((mt[i - 1] ^ (mt[i - 1] >> 30)) * 1566083941u)) - (uint32_t)i;
    ++i;
    if (i >= MT_N) { mt[0] = mt[MT_N - 1]; i = 1; }
  }
  mt[0] = 0x80000000u;
  g->pos = MT_N;
}

// ---- device-side generator working on a shared-memory copy; every lane of the warp calls these together
struct DevMt {
  uint32_t* mt;  // shared
  int pos;
  int has_gauss;
  double gauss;
  int guard;   // guard-band hits of the draws made through this generator
  int regens;  // block regenerations since construction
};

__device__ __forceinline__ uint32_t mt_twist_word(uint32_t cur, uint32_t nxt, uint32_t far) {
  uint32_t y = (cur & 0x80000000u) | (nxt & 0x7fffffffu);
  return far ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
}

// regenerate the 624 words; three dependency-free phases (new[k] needs old[k], old[k+1] and old/new[k+397 mod 624])
__device__ void mt_regenerate(uint32_t* mt) {
  const int lane = threadIdx.x & 31;
  __syncwarp();
  const uint32_t first = mt[0];
  // phase 1: k in [0, 227): uses old[k+397]
  for (int k0 = 0; k0 < MT_N - MT_M; k0 += 32) {
    const int k = k0 + lane;
    uint32_t v = 0;
    if (k < MT_N - MT_M) v = mt_twist_word(mt[k], mt[k + 1], mt[k + MT_M]);
    __syncwarp();
    if (k < MT_N - MT_M) mt[k] = v;
    __syncwarp();
  }
  // phase 2 and 3: k in [227, 623): uses new[k-227]; chunks of <= 227 are independent of themselves
  for (int k0 = MT_N - MT_M; k0 < MT_N - 1; k0 += 32) {
    const int k = k0 + lane;
    uint32_t v = 0;
    if (k < MT_N - 1) v = mt_twist_word(mt[k], mt[k + 1], mt[k - (MT_N - MT_M)]);
    __syncwarp();
    if (k < MT_N - 1) mt[k] = v;
    __syncwarp();
  }
  if (lane == 0) {
    (void)first;
    mt[MT_N - 1] = mt_twist_word(mt[MT_N - 1], mt[0], mt[MT_M - 1]);
  }
  __syncwarp();
}

__device__ __forceinline__ uint32_t mt_u32(DevMt& g) {
  if (g.pos >= MT_N) {  // warp-uniform
    mt_regenerate(g.mt);
    g.pos = 0;
    ++g.regens;
  }
  uint32_t y = g.mt[g.pos++];
  y ^= y >> 11;
  y ^= (y << 7) & 0x9d2c5680u;
  y ^= (y << 15) & 0xefc60000u;
  y ^= y >> 18;
  return y;
}

__device__ __forceinline__ double mt_double(DevMt& g) {
  const uint32_t a = mt_u32(g) >> 5, b = mt_u32(g) >> 6;
  return ((double)a * 67108864.0 + (double)b) / 9007199254740992.0;
}

__device__ __forceinline__ uint32_t py_randbelow(DevMt& g, uint32_t n) {
  const int k = 32 - __clz(n);
  uint32_t r = mt_u32(g) >> (32 - k);
  while (r >= n) r = mt_u32(g) >> (32 - k);
  return r;
}

__device__ __forceinline__ uint32_t np_choice_index(DevMt& g, uint32_t m) {
  const uint32_t rng = m - 1;
  if (rng == 0) return 0;
  uint32_t mask = rng;
  mask |= mask >> 1; mask |= mask >> 2; mask |= mask >> 4; mask |= mask >> 8; mask |= mask >> 16;
  for (;;) {
    const uint32_t v = mt_u32(g) & mask;
    if (v <= rng) return v;
  }
}

// ---- libm guard band ------------------------------------------------------------------------------------------------
// The legacy samplers decide accept/reject by comparing fp64 expressions that contain log().  glibc's log (what the
// reference runs on) is correctly rounded in all but a vanishing fraction of arguments; CUDA's is within 1 ulp of that.
// A comparison whose two sides are further apart than any such difference decides the same way on both; one that falls
// inside the band below is decided again with logarithms computed in double-double arithmetic and rounded once
// (cr_log), i.e. with the values glibc returns.  Everything else in the expressions is +, -, *, /, sqrt, which are
// IEEE-exact on both sides.  (The Gaussian x entering the test carries its own log; a 1-ulp difference there moves both
// sides by ~1e-16 relative, four thousand times less than the band, so it cannot change which side of the band a
// comparison falls on - it can only matter inside a band hit, which the counters report.)
struct dd { double hi, lo; };
__device__ __forceinline__ dd dd_two_sum(double a, double b) {
  const double s = __dadd_rn(a, b), bb = __dadd_rn(s, -a);
  return dd{s, __dadd_rn(__dadd_rn(a, -__dadd_rn(s, -bb)), __dadd_rn(b, -bb))};
}
__device__ __forceinline__ dd dd_quick(double a, double b) {
  const double s = __dadd_rn(a, b);
  return dd{s, __dadd_rn(b, -__dadd_rn(s, -a))};
}
__device__ __forceinline__ dd dd_prod(double a, double b) {
  const double p = __dmul_rn(a, b);
  return dd{p, __fma_rn(a, b, -p)};
}
__device__ __forceinline__ dd dd_add(dd a, dd b) {
  dd s = dd_two_sum(a.hi, b.hi);
  const dd t = dd_two_sum(a.lo, b.lo);
  s.lo = __dadd_rn(s.lo, t.hi);
  s = dd_quick(s.hi, s.lo);
  s.lo = __dadd_rn(s.lo, t.lo);
  return dd_quick(s.hi, s.lo);
}
__device__ __forceinline__ dd dd_mul(dd a, dd b) {
  dd p = dd_prod(a.hi, b.hi);
  p.lo = __dadd_rn(p.lo, __dadd_rn(__dmul_rn(a.hi, b.lo), __dmul_rn(a.lo, b.hi)));
  return dd_quick(p.hi, p.lo);
}
__device__ __forceinline__ dd dd_neg(dd a) { return dd{-a.hi, -a.lo}; }
__device__ __forceinline__ dd dd_div(dd a, dd b) {
  const double q1 = a.hi / b.hi;
  dd r = dd_add(a, dd_neg(dd_mul(b, dd{q1, 0.0})));
  const double q2 = r.hi / b.hi;
  r = dd_add(r, dd_neg(dd_mul(b, dd{q2, 0.0})));
  const double q3 = r.hi / b.hi;
  return dd_add(dd_quick(q1, q2), dd{q3, 0.0});
}
// log(x), x > 0 finite, to ~2^-100 and rounded once to binary64: x = 2^e m with m in [sqrt(1/2), sqrt(2)),
// log m = 2 atanh(s), s = (m - 1)/(m + 1), |s| <= 0.1716: 22 terms of the odd series in double-double
__device__ __noinline__ double cr_log(double x) {
  int e;
  double m = frexp(x, &e);  // m in [0.5, 1)
  if (m < 0.70710678118654752440) { m = __dmul_rn(m, 2.0); --e; }
  const dd num = dd{__dadd_rn(m, -1.0), 0.0};  // exact (Sterbenz)
  const dd den = dd_two_sum(m, 1.0);
  const dd s = dd_div(num, den);
  const dd s2 = dd_mul(s, s);
  dd acc = dd_div(dd{1.0, 0.0}, dd{43.0, 0.0});
  for (int k = 20; k >= 0; --k) acc = dd_add(dd_mul(acc, s2), dd_div(dd{1.0, 0.0}, dd{(double)(2 * k + 1), 0.0}));
  dd r = dd_mul(s, acc);
  r = dd{__dmul_rn(r.hi, 2.0), __dmul_rn(r.lo, 2.0)};
  const dd ln2 = dd{0.69314718055994530942, 2.3190468138462995584e-17};
  r = dd_add(dd_mul(ln2, dd{(double)e, 0.0}), r);
  return __dadd_rn(r.hi, r.lo);
}

// second acceptance test of the Marsaglia-Tsang gamma loop: log(U) < 0.5*X*X + b*(1 - V + log(V)), evaluated without
// fused multiply-add in the operation order of the C source of NumPy's legacy generator
__device__ __forceinline__ bool mt_accept(double u, double x, double v, double b, int& guard) {
  const double lu = log(u), lv = log(v);
  const double hx = __dmul_rn(__dmul_rn(0.5, x), x), omv = __dadd_rn(1.0, -v);
  const double rhs = __dadd_rn(hx, __dmul_rn(b, __dadd_rn(omv, lv)));
  // the band scales with the terms that cancel inside rhs (b*(1-v) against b*log(v) for large shapes)
  const double band = 0x1p-40 * (fabs(lu) + fabs(hx) + fabs(__dmul_rn(b, omv)) + fabs(__dmul_rn(b, lv)));
  if (fabs(__dadd_rn(lu, -rhs)) >= band) return lu < rhs;
  ++guard;
  const double luc = cr_log(u), lvc = cr_log(v);
  return luc < __dadd_rn(hx, __dmul_rn(b, __dadd_rn(omv, lvc)));
}
// Johnk's test X + Y <= 1 with X = U^(1/a): the bandit's shapes are 1 + an integer count, so this branch only ever runs
// with a = b = 1, where X = U and Y = V exactly (no pow at all).  Other shapes <= 1 cannot come from the class API; a
// near-tie there is counted (it is decided with pow as it stands).
__device__ __forceinline__ double johnk_pow(double u, double a) { return a == 1.0 ? u : pow(u, 1.0 / a); }
__device__ __forceinline__ void johnk_guard(double xpy, double a, double b, int& guard) {
  if ((a != 1.0 || b != 1.0) && fabs(__dadd_rn(xpy, -1.0)) < 0x1p-40) ++guard;
}

__device__ double np_gauss(DevMt& g) {
  if (g.has_gauss) {
    const double t = g.gauss;
    g.has_gauss = 0;
    g.gauss = 0.0;
    return t;
  }
  double x1, x2, r2;
  do {
    x1 = 2.0 * mt_double(g) - 1.0;
    x2 = 2.0 * mt_double(g) - 1.0;
    r2 = __dadd_rn(__dmul_rn(x1, x1), __dmul_rn(x2, x2));
  } while (r2 >= 1.0 || r2 == 0.0);
  const double f = sqrt(__dmul_rn(-2.0, log(r2)) / r2);
  g.gauss = __dmul_rn(f, x1);
  g.has_gauss = 1;
  return __dmul_rn(f, x2);
}

__device__ double np_std_gamma(DevMt& g, double shape) {
  if (shape == 1.0) return -log(1.0 - mt_double(g));
  if (shape == 0.0) return 0.0;
  if (shape < 1.0) {
    for (;;) {
      const double u = mt_double(g), v = -log(1.0 - mt_double(g));
      if (u <= 1.0 - shape) {
        const double x = pow(u, 1.0 / shape);
        if (x <= v) return x;
      } else {
        const double y = -log((1 - u) / shape);
        const double x = pow(__dadd_rn(1.0 - shape, __dmul_rn(shape, y)), 1.0 / shape);
        if (x <= __dadd_rn(v, y)) return x;
      }
    }
  }
  const double b = shape - 1.0 / 3.0, c = 1.0 / sqrt(9 * b);
  for (;;) {
    double x, v;
    do {
      x = np_gauss(g);
      v = __dadd_rn(1.0, __dmul_rn(c, x));
    } while (v <= 0.0);
    v = __dmul_rn(__dmul_rn(v, v), v);
    const double u = mt_double(g);
    const double x2 = __dmul_rn(x, x);
    if (u < __dadd_rn(1.0, -__dmul_rn(__dmul_rn(0.0331, x2), x2))) return __dmul_rn(b, v);
    if (mt_accept(u, x, v, b, g.guard)) return __dmul_rn(b, v);
  }
}

// rk_beta draw by draw: the variate is num / den
__device__ void np_beta_parts(DevMt& g, double a, double b, double* num, double* den) {
  *den = 1.0;
  if (a <= 1.0 && b <= 1.0) {
    for (;;) {
      const double u = mt_double(g), v = mt_double(g);
      const double x = johnk_pow(u, a), y = johnk_pow(v, b);
      const double xpy = __dadd_rn(x, y);
      johnk_guard(xpy, a, b, g.guard);
      if (xpy <= 1.0 && __dadd_rn(u, v) > 0.0) {
        if (xpy > 0) { *num = x; *den = xpy; return; }
        double lx = log(u) / a, ly = log(v) / b;
        const double lm = lx > ly ? lx : ly;
        lx -= lm; ly -= lm;
        *num = exp(lx - log(__dadd_rn(exp(lx), exp(ly))));
        return;
      }
    }
  }
  const double ga = np_std_gamma(g, a), gb = np_std_gamma(g, b);
  *num = ga;
  *den = __dadd_rn(ga, gb);
}

// ---- Thompson sampling: the NumPy stream through a look-ahead window ------------------------------------------
// A route draws n beta variates = 2n gamma variates one after the other from one MT19937 stream; evaluated draw by
// draw (np_beta above) that is ~600 dependent fp64 instructions per gamma, 50 us per route at 32 arms - the speed of the
// CPU reference.  The stream positions a draw will read are not known in advance (rejection sampling), but everything
// expensive depends only on the POSITION: the uniform at double-index i, and the polar Gaussian attempt that starts at
// i (two uniforms, a log, a square root, a division).  So the 32 lanes fill tables for the next kWinD positions in
// parallel, and the sequential walk over the arms (same branches, same arithmetic, same order of consumption as
// rk_beta / rk_standard_gamma / rk_gauss of NumPy's legacy generator) only looks values up.
constexpr int kWinD = 304;  // doubles per window: 2 * (kWinD + 2) words span at most two 624-word blocks

struct NpWindow {
  uint32_t *blk0, *blk1;  // shared: two 624-word blocks; `cur` says which one is the committed block
  int cur, next_valid;
  int pos;           // word position in blk[cur] where the window starts (0 .. 624)
  int head;          // doubles consumed from the window
  int has_gauss;
  double gauss;
  int fail;          // a single beta draw ran past the window (probability ~1e-20): that arm is redrawn draw by draw
  double *U, *gret, *gcache, *E;
  uint8_t* ok;
  short* nok;        // nok[i] = first position j >= i, j = i mod 2, whose polar attempt succeeds (kWinD: none in reach)
  int have_E;
  int guard;         // guard-band hits of the draws made through this (copy of the) window
};

__device__ __forceinline__ uint32_t* win_blk(const NpWindow& w, int which) { return which ? w.blk1 : w.blk0; }

__device__ __forceinline__ uint32_t mt_temper(uint32_t y) {
  y ^= y >> 11;
  y ^= (y << 7) & 0x9d2c5680u;
  y ^= (y << 15) & 0xefc60000u;
  y ^= y >> 18;
  return y;
}

// The tables are filled by the whole CTA (kRouterThreads threads): warp 0 runs the bandit and posts a command, the
// other warps wait on a named barrier, fill their share and go back to waiting.
constexpr int kRouterThreads = 256;
struct WinCmd { int p, cur, want_E, exit; };

__device__ __forceinline__ void router_bar() { asm volatile("bar.sync 1, %0;" ::"n"(kRouterThreads) : "memory"); }

__device__ __forceinline__ void win_fill_uniforms(int tid, int p, const uint32_t* b0, const uint32_t* b1, double* U) {
  for (int i = tid; i < kWinD + 2; i += kRouterThreads) {
    const int k = p + 2 * i;
    const uint32_t wa = mt_temper(k < MT_N ? b0[k] : b1[k - MT_N]);
    const uint32_t wb = mt_temper(k + 1 < MT_N ? b0[k + 1] : b1[k + 1 - MT_N]);
    U[i] = ((double)(wa >> 5) * 67108864.0 + (double)(wb >> 6)) / 9007199254740992.0;
  }
}
__device__ __forceinline__ void win_fill_gauss(int tid, bool want_E, const double* U, double* gret, double* gcache,
                                               double* E, uint8_t* ok) {
  for (int i = tid; i < kWinD; i += kRouterThreads) {
    const double x1 = 2.0 * U[i] - 1.0, x2 = 2.0 * U[i + 1] - 1.0;
    const double r2 = __dadd_rn(__dmul_rn(x1, x1), __dmul_rn(x2, x2));
    const bool good = !(r2 >= 1.0 || r2 == 0.0);
    ok[i] = good;
    if (good) {
      const double f = sqrt(__dmul_rn(-2.0, log(r2)) / r2);
      gcache[i] = __dmul_rn(f, x1);
      gret[i] = __dmul_rn(f, x2);
    }
    if (want_E) E[i] = -log(1.0 - U[i]);
  }
}

__device__ __forceinline__ void win_fill_next_ok(int tid, const uint8_t* ok, short* nok) {
  for (int i = tid; i < kWinD; i += kRouterThreads) {
    int j = i;
    while (j < kWinD && !ok[j]) j += 2;
    nok[i] = (short)(j < kWinD ? j : kWinD);
  }
}

// move the committed position to the first unconsumed word and fill the tables from there (called by warp 0)
__device__ __forceinline__ void np_window_build(NpWindow& w, bool want_E, WinCmd* cmd) {
  const int lane = threadIdx.x & 31;
  int p = w.pos + 2 * w.head;
  if (p > MT_N) {  // past the end of the committed block: the next block becomes the committed one
    if (!w.next_valid) {
      for (int i = lane; i < MT_N; i += 32) win_blk(w, w.cur ^ 1)[i] = win_blk(w, w.cur)[i];
      mt_regenerate(win_blk(w, w.cur ^ 1));
    }
    w.cur ^= 1;
    w.next_valid = 0;
    p -= MT_N;
  }
  w.pos = p;
  w.head = 0;
  if (p + 2 * (kWinD + 2) > MT_N && !w.next_valid) {
    __syncwarp();
    for (int i = lane; i < MT_N; i += 32) win_blk(w, w.cur ^ 1)[i] = win_blk(w, w.cur)[i];
    mt_regenerate(win_blk(w, w.cur ^ 1));
    w.next_valid = 1;
  }
  __syncwarp();
  if (lane == 0) *cmd = WinCmd{p, w.cur, want_E ? 1 : 0, 0};
  __syncwarp();
  router_bar();  // releases the helper warps with the command
  win_fill_uniforms(lane, p, win_blk(w, w.cur), win_blk(w, w.cur ^ 1), w.U);
  router_bar();
  win_fill_gauss(lane, want_E, w.U, w.gret, w.gcache, w.E, w.ok);
  router_bar();
  win_fill_next_ok(lane, w.ok, w.nok);
  router_bar();
  w.have_E = want_E;
}

__device__ __forceinline__ double win_double(NpWindow& w) {
  if (w.head >= kWinD) { w.fail = 1; w.head = kWinD - 1; }
  return w.U[w.head++];
}
__device__ __forceinline__ double win_exponential(NpWindow& w) {
  if (w.head >= kWinD) { w.fail = 1; w.head = kWinD - 1; }
  const int i = w.head++;
  return w.have_E ? w.E[i] : -log(1.0 - w.U[i]);
}
__device__ __forceinline__ double win_gauss(NpWindow& w) {
  if (w.has_gauss) {
    const double t = w.gauss;
    w.has_gauss = 0;
    w.gauss = 0.0;
    return t;
  }
  while (w.head < kWinD - 1 && !w.ok[w.head]) w.head += 2;
  if (w.head >= kWinD - 1) { w.fail = 1; w.head = kWinD - 2; return 0.0; }
  const int i = w.head;
  w.head += 2;
  w.gauss = w.gcache[i];
  w.has_gauss = 1;
  return w.gret[i];
}

// rk_standard_gamma of the legacy generator, reading the stream through the window (same arithmetic as np_std_gamma)
// c = 1/sqrt(9 (shape - 1/3)) of the Marsaglia-Tsang loop, tabulated per arm when c_tab >= 0 (it only changes when the
// arm receives feedback); a negative c_tab means "compute here"
__device__ __forceinline__ double gamma_c(double shape) { return shape > 1.0 ? 1.0 / sqrt(9 * (shape - 1.0 / 3.0)) : 0.0; }

__device__ __forceinline__ double win_std_gamma(NpWindow& w, double shape, double c_tab) {
  if (shape == 1.0) return win_exponential(w);
  if (shape == 0.0) return 0.0;
  if (shape < 1.0) {
    for (;;) {
      const double u = win_double(w), v = win_exponential(w);
      if (w.fail) return 0.0;
      if (u <= 1.0 - shape) {
        const double x = pow(u, 1.0 / shape);
        if (x <= v) return x;
      } else {
        const double y = -log((1 - u) / shape);
        const double x = pow(__dadd_rn(1.0 - shape, __dmul_rn(shape, y)), 1.0 / shape);
        if (x <= __dadd_rn(v, y)) return x;
      }
    }
  }
  const double b = shape - 1.0 / 3.0, c = c_tab >= 0.0 ? c_tab : 1.0 / sqrt(9 * b);
  for (;;) {
    double x, v;
    do {
      x = win_gauss(w);
      v = __dadd_rn(1.0, __dmul_rn(c, x));
    } while (v <= 0.0 && !w.fail);
    v = __dmul_rn(__dmul_rn(v, v), v);
    const double u = win_double(w);
    if (w.fail) return 0.0;
    const double x2 = __dmul_rn(x, x);
    if (u < __dadd_rn(1.0, -__dmul_rn(__dmul_rn(0.0331, x2), x2))) return __dmul_rn(b, v);
    if (mt_accept(u, x, v, b, w.guard)) return __dmul_rn(b, v);
  }
}

// rk_beta: the variate is num / den (the division is left to the caller, which does all arms at once)
__device__ __forceinline__ void win_beta(NpWindow& w, double a, double b, double ca, double cb, double* num, double* den) {
  *den = 1.0;
  *num = 0.0;
  if (a <= 1.0 && b <= 1.0) {
    for (;;) {
      const double u = win_double(w), v = win_double(w);
      if (w.fail) return;
      const double x = johnk_pow(u, a), y = johnk_pow(v, b);
      const double xpy = __dadd_rn(x, y);
      johnk_guard(xpy, a, b, w.guard);
      if (xpy <= 1.0 && __dadd_rn(u, v) > 0.0) {
        if (xpy > 0) { *num = x; *den = xpy; return; }
        double lx = log(u) / a, ly = log(v) / b;
        const double lm = lx > ly ? lx : ly;
        lx -= lm; ly -= lm;
        *num = exp(lx - log(__dadd_rn(exp(lx), exp(ly))));
        return;
      }
    }
  }
  const double ga = win_std_gamma(w, a, ca), gb = win_std_gamma(w, b, cb);
  *num = ga;
  *den = __dadd_rn(ga, gb);
}

// One beta draw taken draw by draw from the committed stream position, for the arm whose draw outran the look-ahead
// window (or on request, SB_ROUTER_OPT_DEBUG_FALLBACK_PERIOD).  `w` holds the state before that arm; on return it is
// the state after it, and the tables must be rebuilt.  Called by warp 0 with all lanes in step.
__device__ void win_fallback_beta(NpWindow& w, double a, double b, double* num, double* den) {
  const int lane = threadIdx.x & 31;
  int p = w.pos + 2 * w.head;
  if (p >= MT_N) {  // the committed position lies in the next block
    if (!w.next_valid) {
      for (int i = lane; i < MT_N; i += 32) win_blk(w, w.cur ^ 1)[i] = win_blk(w, w.cur)[i];
      mt_regenerate(win_blk(w, w.cur ^ 1));
    }
    w.cur ^= 1;
    w.next_valid = 0;
    p -= MT_N;
  }
  __syncwarp();
  DevMt g{win_blk(w, w.cur), p, w.has_gauss, w.gauss, 0, 0};
  np_beta_parts(g, a, b, num, den);
  if (g.regens) w.next_valid = 0;  // the block was regenerated in place: the other one no longer is its successor
  w.pos = g.pos;
  w.head = 0;
  w.has_gauss = g.has_gauss;
  w.gauss = g.gauss;
  w.guard += g.guard;
  w.fail = 0;
  __syncwarp();
}

// T calls.  KIND 0 = epsilon-greedy, 1 = Thompson (a template parameter so that the epsilon-greedy instance does not carry
// the Thompson window code and its registers).  reward_mode -1: route only.
template <int KIND>
__global__ void __launch_bounds__(kRouterThreads)
router_run_kernel(RouterState* __restrict__ st, int n, double epsilon, long long T,
                  const long long* __restrict__ rows, const float* __restrict__ reward,
                  const float* __restrict__ arm_p, int reward_mode, int* __restrict__ decisions, int forced_arm,
                  double forced_reward, long long forced_rows, int dbg_fallback_period, int pre_arm,
                  double pre_reward, long long pre_rows) {
  constexpr int kind = KIND;
  __shared__ uint32_t s_py[MT_N], s_np[MT_N], s_np2[MT_N];
  __shared__ double s_U[kWinD + 2], s_gret[kWinD], s_gcache[kWinD], s_E[kWinD];
  __shared__ uint8_t s_ok[kWinD];
  __shared__ short s_nok[kWinD];
  constexpr int kTabArms = 128;  // arms with tabulated gamma constants and a parallel division (beyond: inline)
  __shared__ double s_ca[kTabArms], s_cb[kTabArms], s_num[kTabArms], s_den[kTabArms];
  __shared__ uint8_t s_kind[kTabArms];  // bit 0: a == 1, bit 1: b == 1 (unit shapes draw an exponential, no Gaussian)
  __shared__ long long s_succ[kMaxArms], s_tries[kMaxArms];
  __shared__ double s_val[kMaxArms];
  __shared__ WinCmd s_cmd;
  if (threadIdx.x >= 32) {
    // helper warps: table fills on command until warp 0 says exit
    for (;;) {
      router_bar();
      const WinCmd c = s_cmd;
      if (c.exit) return;
      const uint32_t* b0 = c.cur ? s_np2 : s_np;
      const uint32_t* b1 = c.cur ? s_np : s_np2;
      win_fill_uniforms(threadIdx.x, c.p, b0, b1, s_U);
      router_bar();
      win_fill_gauss(threadIdx.x, c.want_E != 0, s_U, s_gret, s_gcache, s_E, s_ok);
      router_bar();
      win_fill_next_ok(threadIdx.x, s_ok, s_nok);
      router_bar();
    }
  }
  const int lane = threadIdx.x;
  for (int i = lane; i < MT_N; i += 32) { s_py[i] = st->py.mt[i]; s_np[i] = st->np.mt[i]; }
  for (int j = lane; j < n; j += 32) {
    s_succ[j] = st->succ[j];
    s_tries[j] = st->tries[j];
    s_val[j] = st->tries[j] ? (double)st->succ[j] / (double)st->tries[j] : 0.0;
  }
  DevMt py{s_py, st->py.pos, 0, 0.0, 0, 0};
  DevMt np{s_np, st->np.pos, st->np.has_gauss, st->np.gauss, 0, 0};
  long long n_fallback = 0, n_draws = 0;
  int best = st->best;
  const bool tab = kind == 1 && n <= kTabArms;
  if (tab)
    for (int j = lane; j < n; j += 32) {
      s_ca[j] = gamma_c((double)(1 + s_succ[j]));
      s_cb[j] = gamma_c((double)(1 + s_tries[j] - s_succ[j]));
      s_kind[j] = (uint8_t)((s_succ[j] == 0 ? 1 : 0) | (s_tries[j] == s_succ[j] ? 2 : 0));
    }
  __syncwarp();
  // Thompson: the NumPy stream is read through a window (committed position = np.pos at entry)
  NpWindow win{s_np, s_np2, 0, 0, np.pos, 0, np.has_gauss, np.gauss, 0, s_U, s_gret, s_gcache, s_E, s_ok, s_nok, 0, 0};
  const bool windowed = kind == 1 && forced_arm < 0;
  // room one beta draw may need: two gammas, each a few doubles per trial; a refill is forced below this margin
  constexpr int kArmMargin = 48;
  int win_built = 0;
  // pre_arm >= 0: one feedback (for a call routed by an earlier launch) is applied before the T calls of this launch -
  // send_feedback followed by the next route() of the class API costs one launch instead of two
  for (long long t = pre_arm >= 0 ? -1 : 0; t < T; ++t) {
    const bool pre = t < 0;
    int arm;
    if (pre) {
      arm = pre_arm;
    } else if (forced_arm >= 0) {
      arm = forced_arm;  // send_feedback for a call routed earlier
    } else if (kind == 0) {
      const double x = mt_double(py);  // EpsilonGreedy.py:90
      if (x > epsilon) arm = best;
      else {
        const int r = (int)py_randbelow(py, (uint32_t)(n - 1));  // L91-94
        arm = r < best ? r : r + 1;
      }
    } else {
      double top = 0.0;
      arm = 0;
      // unit shapes draw an exponential: tabulate -log(1 - U) with the window when some arm still has one
      bool unit = false;
      for (int j = lane; j < n; j += 32) unit |= s_succ[j] == 0 || s_tries[j] == s_succ[j];
      unit = __any_sync(0xffffffffu, unit);
      if (tab) {
        // Arms as lanes.  The draws of one route read the stream one after the other, but nearly always in a pattern
        // that the tables give away: a gamma with shape > 1 takes the cached Gaussian or else the first successful polar
        // pair at or after its position (nok[]) and one uniform (the squeeze test accepts ~97 %), a unit shape takes
        // one uniform, two unit shapes (Johnk) take pairs of uniforms until their sum is <= 1.  The warp follows that
        // pattern arm by arm (cheap, table look-ups only) and lane l latches the state PREDICTED for arm done + l;
        // each lane then draws its beta from there on a private copy of the window state, and the warp accepts the
        // longest prefix of lanes whose assumed start state equals the actual end state of the lane before (position,
        // cached-Gaussian flag and value).  Everything accepted is exactly what the sequential walk would have drawn.
        int done = 0;
        while (done < n) {
          if (!win_built || kWinD - win.head < kArmMargin) {
            np_window_build(win, unit, &s_cmd);
            win_built = 1;
          }
          // test knob: draws until the next one that is forced through the fallback (0 = this one)
          const int dbg_left = dbg_fallback_period > 0
                                   ? dbg_fallback_period - 1 - (int)((n_draws + done) % dbg_fallback_period) : 1 << 30;
          if (dbg_left == 0) {
            const double a = (double)(1 + s_succ[done]), b = (double)(1 + s_tries[done] - s_succ[done]);
            double num, den;
            win_fallback_beta(win, a, b, &num, &den);
            if (lane == 0) { s_num[done] = num; s_den[done] = den; }
            ++n_fallback; ++done; win_built = 0;
            continue;
          }
          const int D = min(min(n - done, 32), dbg_left);
          int p = win.head, f = win.has_gauss, gq = -1;   // lane 0: the actual state
          {
            int cp = p, cf = f, cq = -1;  // the chase, identical in every lane
            for (int it = 1; it < D; ++it) {
              const int kd = s_kind[done + it - 1];
              if (kd == 3) {
                int tries = 0;
                while (cp + 1 < kWinD && tries < 16 && !(__dadd_rn(s_U[cp], s_U[cp + 1]) <= 1.0)) { cp += 2; ++tries; }
                cp += 2;
              } else if (cf) {
                if (kd == 0) {
                  const int q = cp + 1 < kWinD ? (int)s_nok[cp + 1] : kWinD;
                  cp = q + 3;
                  cq = q;
                } else {
                  cp += 2;
                  cf = 0;
                }
              } else {
                const int at = kd == 1 ? cp + 1 : cp;
                const int q = at < kWinD ? (int)s_nok[at] : kWinD;
                cp = kd == 1 ? q + 3 : q + 4;
                if (kd != 0) { cf = 1; cq = q; }
              }
              if (cp > kWinD) cp = kWinD;
              if (lane == it) { p = cp; f = cf; gq = cq; }
            }
          }
          bool valid = lane < D && p <= kWinD - kArmMargin;
          double g_in = win.gauss;
          if (lane > 0) {
            g_in = 0.0;
            if (f) {  // the value the arms before leave in the cache: the other half of the last polar pair
              valid = valid && gq >= 0 && gq < kWinD && s_ok[gq];
              if (valid) g_in = s_gcache[gq];
            }
          }
          const int flag0 = f;
          NpWindow lw = win;
          lw.head = p;
          lw.has_gauss = flag0;
          lw.gauss = g_in;
          lw.fail = 0;
          lw.guard = 0;
          const int j = done + lane;
          double num = 0.0, den = 1.0;
          if (valid) {
            const double a = (double)(1 + s_succ[j]), b = (double)(1 + s_tries[j] - s_succ[j]);
            win_beta(lw, a, b, s_ca[j], s_cb[j], &num, &den);
          }
          const int e_head = lw.head, e_flag = lw.has_gauss;
          const long long e_bits = __double_as_longlong(lw.gauss);
          const int pe_head = __shfl_up_sync(0xffffffffu, e_head, 1), pe_flag = __shfl_up_sync(0xffffffffu, e_flag, 1);
          const long long pe_bits = __shfl_up_sync(0xffffffffu, e_bits, 1);
          const bool okl = valid && !lw.fail &&
                           (lane == 0 || (pe_head == p && pe_flag == flag0 && (!flag0 || pe_bits == __double_as_longlong(g_in))));
          const unsigned bal = __ballot_sync(0xffffffffu, okl);
          const int k = bal == 0xffffffffu ? 32 : __ffs((int)~bal) - 1;  // accepted lanes: 0 .. k-1
          if (k == 0) {
            // the first lane itself ran out of window: that arm is drawn draw by draw from the committed position
            const double a = (double)(1 + s_succ[done]), b = (double)(1 + s_tries[done] - s_succ[done]);
            double fnum, fden;
            win_fallback_beta(win, a, b, &fnum, &fden);
            if (lane == 0) { s_num[done] = fnum; s_den[done] = fden; }
            ++n_fallback; ++done; win_built = 0;
            continue;
          }
          if (lane < k) { s_num[j] = num; s_den[j] = den; }
          {
            int gh = lane < k ? lw.guard : 0;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) gh += __shfl_xor_sync(0xffffffffu, gh, o);
            win.guard += gh;
          }
          win.head = __shfl_sync(0xffffffffu, e_head, k - 1);
          win.has_gauss = __shfl_sync(0xffffffffu, e_flag, k - 1);
          win.gauss = __longlong_as_double(__shfl_sync(0xffffffffu, e_bits, k - 1));
          done += k;
        }
      } else {
        for (int j = 0; j < n; ++j) {  // ThompsonSampling.py:81-82, arm order; np.argmax = first maximum
          if (!win_built || kWinD - win.head < kArmMargin) {
            np_window_build(win, unit, &s_cmd);
            win_built = 1;
          }
          const double a = (double)(1 + s_succ[j]), b = (double)(1 + s_tries[j] - s_succ[j]);
          double num, den;
          const int h0 = win.head, f0 = win.has_gauss, g0 = win.guard;
          const double c0 = win.gauss;
          const bool force = dbg_fallback_period > 0 && (n_draws + j) % dbg_fallback_period == dbg_fallback_period - 1;
          if (!force) win_beta(win, a, b, -1.0, -1.0, &num, &den);
          if (force || win.fail) {
            win.head = h0; win.has_gauss = f0; win.gauss = c0; win.guard = g0;
            win_fallback_beta(win, a, b, &num, &den);
            ++n_fallback; win_built = 0;
          }
          const double v = num / den;
          if (lane == 0 && t == T - 1) st->last_values[j] = v;
          if (j == 0 || v > top) { top = v; arm = j; }
        }
      }
      n_draws += n;
      if (tab) {
        // the n divisions in parallel, then the same first-maximum scan
        __syncwarp();
        for (int j = lane; j < n; j += 32) s_num[j] = s_num[j] / s_den[j];
        __syncwarp();
        for (int j = 0; j < n; ++j) {
          const double v = s_num[j];
          if (j == 0 || v > top) { top = v; arm = j; }
        }
        if (t == T - 1)
          for (int j = lane; j < n; j += 32) st->last_values[j] = s_num[j];
      }
    }
    if (lane == 0 && forced_arm < 0 && !pre) decisions[t] = arm;
    if (reward_mode < 0 && !pre) continue;
    const long long nrows = pre ? pre_rows : (forced_arm >= 0 ? forced_rows : (rows ? rows[t] : 1));
    double r;
    if (pre) r = pre_reward;
    else if (forced_arm >= 0) r = forced_reward;
    else if (reward_mode == 0) r = (double)reward[t];
    else if (reward_mode == 1) r = (double)reward[t * n + arm];
    else r = reward[t] < arm_p[arm] ? 1.0 : 0.0;
    const long long ns = (long long)__dmul_rn(r, (double)nrows);  // int(reward * n_predictions), truncation
    __syncwarp();
    if (lane == 0) {
      s_succ[arm] += ns;
      s_tries[arm] += nrows;
      if (kind == 0) s_val[arm] = (double)s_succ[arm] / (double)s_tries[arm];  // L118-119
      if (tab) {
        s_ca[arm] = gamma_c((double)(1 + s_succ[arm]));
        s_cb[arm] = gamma_c((double)(1 + s_tries[arm] - s_succ[arm]));
        s_kind[arm] = (uint8_t)((s_succ[arm] == 0 ? 1 : 0) | (s_tries[arm] == s_succ[arm] ? 2 : 0));
      }
    }
    __syncwarp();
    if (kind == 0) {
      // best = np.random.choice(where(values == max(values)))   L124-125
      double top = -INFINITY;
      for (int j = lane; j < n; j += 32) top = fmax(top, s_val[j]);
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) top = fmax(top, __shfl_xor_sync(0xffffffffu, top, o));
      int m = 0;
      for (int j0 = 0; j0 < n; j0 += 32) {
        const int j = j0 + lane;
        m += __popc(__ballot_sync(0xffffffffu, j < n && s_val[j] == top));
      }
      int pick = (int)np_choice_index(np, (uint32_t)m);
      for (int j0 = 0; j0 < n; j0 += 32) {
        const int j = j0 + lane;
        const unsigned bal = __ballot_sync(0xffffffffu, j < n && s_val[j] == top);
        const int cnt = __popc(bal);
        if (pick < cnt) {
          unsigned b2 = bal;
          for (int s = 0; s < pick; ++s) b2 &= b2 - 1;
          best = j0 + __ffs(b2) - 1;
          break;
        }
        pick -= cnt;
      }
    }
  }
  __syncwarp();
  if (blockDim.x > 32) {  // Thompson launches carry helper warps: tell them to leave
    if (lane == 0) s_cmd.exit = 1;
    __syncwarp();
    router_bar();
  }
  const uint32_t* np_blk = s_np;
  if (windowed && (win_built || n_fallback > 0)) {
    // commit the consumed part of the window: the state a draw-by-draw generator would be in
    int p = win.pos + 2 * win.head;
    if (p > MT_N) {
      if (!win.next_valid) {
        for (int i = lane; i < MT_N; i += 32) win_blk(win, win.cur ^ 1)[i] = win_blk(win, win.cur)[i];
        mt_regenerate(win_blk(win, win.cur ^ 1));
      }
      win.cur ^= 1;
      p -= MT_N;
    }
    np.pos = p;
    np.has_gauss = win.has_gauss;
    np.gauss = win.gauss;
    np_blk = win_blk(win, win.cur);
    __syncwarp();
  }
  for (int i = lane; i < MT_N; i += 32) { st->py.mt[i] = s_py[i]; st->np.mt[i] = np_blk[i]; }
  for (int j = lane; j < n; j += 32) { st->succ[j] = s_succ[j]; st->tries[j] = s_tries[j]; }
  if (lane == 0) {
    st->py.pos = py.pos;
    st->np.pos = np.pos;
    st->np.has_gauss = np.has_gauss;
    st->np.gauss = np.gauss;
    st->best = best;
    st->guard_hits += (long long)win.guard + np.guard;
    st->fallbacks += n_fallback;
  }
}

}  // namespace

struct sb_router {
  int kind, n;
  double epsilon;
  RouterState* d_state;
  int dbg_fallback_period;
  int* h_decision;  // pinned, device-visible: the kernel of sb_router_route writes the decision straight to the host
};

extern "C" int32_t sb_router_create(sb_router_t** out, int32_t kind, int32_t n_branches, double epsilon, int64_t seed,
                                    int32_t best_branch) {
  SB_RANGE("sb_router_create");
  SB_REQUIRE(out, SB_ERR_INVALID, "sb_router_create: null handle pointer");
  SB_REQUIRE(kind == SB_ROUTER_EPSILON_GREEDY || kind == SB_ROUTER_THOMPSON, SB_ERR_INVALID, "sb_router_create: bad kind");
  SB_REQUIRE(n_branches > 0 && n_branches <= kMaxArms, SB_ERR_INVALID, "n_branches parameter must be given (1..%d)", kMaxArms);
  SB_REQUIRE(seed > 0 && seed <= 0xffffffffll, SB_ERR_INVALID,
             "sb_router_create: seed must be in 1..2^32-1 (np.random.seed range; a falsy seed means unseeded in the "
             "reference - pick one on the host)");
  SB_REQUIRE(best_branch < n_branches, SB_ERR_INVALID, "sb_router_create: best_branch out of range");
  RouterState* hs = new RouterState();
  memset(hs, 0, sizeof *hs);
  uint32_t key[2];
  int len = 0;
  for (uint64_t a = (uint64_t)seed; a; a >>= 32) key[len++] = (uint32_t)a;
  mt_init_by_array(&hs->py, key, len);       // random.seed(seed)      EpsilonGreedy.py:48
  mt_init_genrand(&hs->np, (uint32_t)seed);  // np.random.seed(seed)   EpsilonGreedy.py:49
  hs->best = 0;
  if (kind == SB_ROUTER_EPSILON_GREEDY) {
    if (best_branch > 0) hs->best = best_branch;  // `if best_branch:` - 0 is falsy (EpsilonGreedy.py:64)
    else {
      // random.choice(range(n)): one _randbelow draw from the python stream, done on the host copy
      MtState* g = &hs->py;
      auto u32 = [&]() {
        if (g->pos >= MT_N) {
          for (int k = 0; k < MT_N; ++k) {
            uint32_t y = (g->mt[k] & 0x80000000u) | (g->mt[(k + 1) % MT_N] & 0x7fffffffu);
            g->mt[k] = g->mt[(k + MT_M) % MT_N] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
          }
          g->pos = 0;
        }
        uint32_t y = g->mt[g->pos++];
        y ^= y >> 11; y ^= (y << 7) & 0x9d2c5680u; y ^= (y << 15) & 0xefc60000u; y ^= y >> 18;
        return y;
      };
      const uint32_t n = (uint32_t)n_branches;
      const int k = 32 - __builtin_clz(n);
      uint32_t r = u32() >> (32 - k);
      while (r >= n) r = u32() >> (32 - k);
      hs->best = (int)r;
    }
  }
  sb_router* h = new sb_router{kind, n_branches, epsilon, nullptr, 0, nullptr};
  cudaError_t e = cudaMalloc(&h->d_state, sizeof(RouterState));
  if (e == cudaSuccess) e = cudaHostAlloc(&h->h_decision, sizeof(int), cudaHostAllocMapped | cudaHostAllocPortable);
  if (e == cudaSuccess) e = cudaMemcpy(h->d_state, hs, sizeof(RouterState), cudaMemcpyHostToDevice);
  delete hs;
  if (e != cudaSuccess) {
    sb::set_error("sb_router_create: %s", cudaGetErrorString(e));
    if (h->d_state) cudaFree(h->d_state);
    if (h->h_decision) cudaFreeHost(h->h_decision);
    delete h;
    return SB_ERR_CUDA;
  }
  *out = h;
  return SB_OK;
}

extern "C" int32_t sb_router_destroy(sb_router_t* h) {
  SB_RANGE("sb_router_destroy");
  if (!h) return SB_OK;
  cudaFree(h->d_state);
  cudaFreeHost(h->h_decision);
  delete h;
  return SB_OK;
}

extern "C" int32_t sb_router_run(sb_router_t* h, int64_t T, const int64_t* d_rows_per_call, const float* d_reward,
                                 const float* d_arm_p, int32_t reward_mode, int32_t* d_decisions, uintptr_t stream) {
  SB_RANGE("sb_router_run");
  SB_REQUIRE(h && T >= 0 && reward_mode >= -1 && reward_mode <= 2, SB_ERR_INVALID, "sb_router_run: bad arguments");
  if (T == 0) return SB_OK;
  SB_REQUIRE(d_decisions, SB_ERR_INVALID, "sb_router_run: null decisions");
  SB_REQUIRE(reward_mode < 0 || d_reward, SB_ERR_INVALID, "sb_router_run: rewards required");
  SB_REQUIRE(reward_mode != SB_REWARD_BERNOULLI || d_arm_p, SB_ERR_INVALID, "sb_router_run: arm probabilities required");
  SB_REQUIRE(!(h->kind == SB_ROUTER_EPSILON_GREEDY && h->n < 2), SB_ERR_INVALID,
             "epsilon-greedy needs >= 2 branches to explore (reference raises IndexError on random.choice([]))");
  if (h->kind == SB_ROUTER_THOMPSON)
    router_run_kernel<1><<<1, kRouterThreads, 0, (cudaStream_t)stream>>>(h->d_state, h->n, h->epsilon, (long long)T,
                                                                        (const long long*)d_rows_per_call, d_reward,
                                                                        d_arm_p, reward_mode, d_decisions, -1, 0.0, 0,
                                                                        h->dbg_fallback_period, -1, 0.0, 0);
  else
    router_run_kernel<0><<<1, 32, 0, (cudaStream_t)stream>>>(h->d_state, h->n, h->epsilon, (long long)T,
                                                            (const long long*)d_rows_per_call, d_reward, d_arm_p,
                                                            reward_mode, d_decisions, -1, 0.0, 0, 0, -1, 0.0, 0);
  SB_LAUNCH_CHECK();
  return SB_OK;
}

extern "C" int32_t sb_router_feedback(sb_router_t* h, int32_t arm, int64_t rows, double reward, uintptr_t stream) {
  SB_RANGE("sb_router_feedback");
  SB_REQUIRE(h && arm >= 0 && arm < h->n, SB_ERR_INVALID, "sb_router_feedback: routing %d outside 0..%d", arm,
             h ? h->n - 1 : 0);
  SB_REQUIRE(rows >= 1, SB_ERR_INVALID, "sb_router_feedback: rows must be >= 1 (reference divides by tries)");
  if (h->kind == SB_ROUTER_THOMPSON)
    router_run_kernel<1><<<1, kRouterThreads, 0, (cudaStream_t)stream>>>(h->d_state, h->n, h->epsilon, 1, nullptr, nullptr,
                                                                        nullptr, 0, nullptr, arm, reward, (long long)rows, 0, -1, 0.0, 0);
  else
    router_run_kernel<0><<<1, 32, 0, (cudaStream_t)stream>>>(h->d_state, h->n, h->epsilon, 1, nullptr, nullptr, nullptr, 0,
                                                            nullptr, arm, reward, (long long)rows, 0, -1, 0.0, 0);
  SB_LAUNCH_CHECK();
  return SB_OK;
}

extern "C" int64_t sb_router_state_len(const sb_router_t* h) {
  return h ? 2 * (int64_t)h->n + 1 + (MT_N + 1) + (MT_N + 3) : 0;
}

extern "C" int32_t sb_router_get_state(sb_router_t* h, int64_t* out, uintptr_t stream) {
  SB_RANGE("sb_router_get_state");
  SB_REQUIRE(h && out, SB_ERR_INVALID, "sb_router_get_state: bad arguments");
  RouterState* hs = new RouterState();
  cudaStream_t st = (cudaStream_t)stream;
  cudaError_t e = cudaMemcpyAsync(hs, h->d_state, sizeof(RouterState), cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  if (e != cudaSuccess) { delete hs; sb::set_error("sb_router_get_state: %s", cudaGetErrorString(e)); return SB_ERR_CUDA; }
  int64_t* p = out;
  for (int j = 0; j < h->n; ++j) *p++ = hs->succ[j];
  for (int j = 0; j < h->n; ++j) *p++ = hs->tries[j];
  *p++ = hs->best;
  for (int i = 0; i < MT_N; ++i) *p++ = hs->py.mt[i];
  *p++ = hs->py.pos;
  for (int i = 0; i < MT_N; ++i) *p++ = hs->np.mt[i];
  *p++ = hs->np.pos;
  *p++ = hs->np.has_gauss;
  long long bits;
  memcpy(&bits, &hs->np.gauss, 8);
  *p++ = bits;
  delete hs;
  return SB_OK;
}

extern "C" int32_t sb_router_set_state(sb_router_t* h, const int64_t* in, uintptr_t stream) {
  SB_RANGE("sb_router_set_state");
  SB_REQUIRE(h && in, SB_ERR_INVALID, "sb_router_set_state: bad arguments");
  RouterState* hs = new RouterState();
  memset(hs, 0, sizeof *hs);
  const int64_t* p = in;
  for (int j = 0; j < h->n; ++j) hs->succ[j] = *p++;
  for (int j = 0; j < h->n; ++j) hs->tries[j] = *p++;
  hs->best = (int)*p++;
  for (int i = 0; i < MT_N; ++i) hs->py.mt[i] = (uint32_t)*p++;
  hs->py.pos = (int)*p++;
  for (int i = 0; i < MT_N; ++i) hs->np.mt[i] = (uint32_t)*p++;
  hs->np.pos = (int)*p++;
  hs->np.has_gauss = (int)*p++;
  long long bits = *p++;
  memcpy(&hs->np.gauss, &bits, 8);
  cudaStream_t st = (cudaStream_t)stream;
  cudaError_t e = cudaMemcpyAsync(h->d_state, hs, sizeof(RouterState), cudaMemcpyHostToDevice, st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  delete hs;
  if (e != cudaSuccess) { sb::set_error("sb_router_set_state: %s", cudaGetErrorString(e)); return SB_ERR_CUDA; }
  return SB_OK;
}

extern "C" int32_t sb_router_set_option(sb_router_t* h, int32_t option, int64_t value) {
  SB_RANGE("sb_router_set_option");
  SB_REQUIRE(h, SB_ERR_INVALID, "sb_router_set_option: null handle");
  SB_REQUIRE(option == SB_ROUTER_OPT_DEBUG_FALLBACK_PERIOD && value >= 0 && value <= 0x7fffffff, SB_ERR_INVALID,
             "sb_router_set_option: unknown option %d or bad value", option);
  h->dbg_fallback_period = (int)value;
  return SB_OK;
}

extern "C" int32_t sb_router_counters(sb_router_t* h, int64_t* out2, uintptr_t stream) {
  SB_RANGE("sb_router_counters");
  SB_REQUIRE(h && out2, SB_ERR_INVALID, "sb_router_counters: bad arguments");
  long long v[2];
  cudaStream_t st = (cudaStream_t)stream;
  SB_CUDA(cudaMemcpyAsync(v, &h->d_state->guard_hits, sizeof v, cudaMemcpyDeviceToHost, st));
  SB_CUDA(cudaStreamSynchronize(st));
  out2[0] = v[0];
  out2[1] = v[1];
  return SB_OK;
}

// route() of the class API: one call, the decision written by the kernel into pinned host memory (no device buffer, no
// copy), then a stream synchronisation.  pre_arm >= 0: the feedback of the previous request is applied first, in the
// same launch.
static int32_t router_route_impl(sb_router_t* h, int32_t pre_arm, int64_t pre_rows, double pre_reward, int32_t* h_out,
                                 uintptr_t stream) {
  SB_REQUIRE(h && h_out, SB_ERR_INVALID, "sb_router_route: bad arguments");
  SB_REQUIRE(!(h->kind == SB_ROUTER_EPSILON_GREEDY && h->n < 2), SB_ERR_INVALID,
             "epsilon-greedy needs >= 2 branches to explore (reference raises IndexError on random.choice([]))");
  SB_REQUIRE(pre_arm < h->n && (pre_arm < 0 || pre_rows >= 1), SB_ERR_INVALID,
             "sb_router_feedback_route: routing %d outside 0..%d or rows < 1", pre_arm, h->n - 1);
  int* d_dec = nullptr;
  SB_CUDA(cudaHostGetDevicePointer((void**)&d_dec, h->h_decision, 0));
  if (h->kind == SB_ROUTER_THOMPSON)
    router_run_kernel<1><<<1, kRouterThreads, 0, (cudaStream_t)stream>>>(h->d_state, h->n, h->epsilon, 1, nullptr, nullptr,
                                                                        nullptr, -1, d_dec, -1, 0.0, 0,
                                                                        h->dbg_fallback_period, pre_arm, pre_reward,
                                                                        (long long)pre_rows);
  else
    router_run_kernel<0><<<1, 32, 0, (cudaStream_t)stream>>>(h->d_state, h->n, h->epsilon, 1, nullptr, nullptr, nullptr, -1,
                                                            d_dec, -1, 0.0, 0, 0, pre_arm, pre_reward,
                                                            (long long)pre_rows);
  SB_LAUNCH_CHECK();
  SB_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
  *h_out = *(volatile int*)h->h_decision;
  return SB_OK;
}

extern "C" int32_t sb_router_route(sb_router_t* h, int32_t* h_out, uintptr_t stream) {
  SB_RANGE("sb_router_route");
  return router_route_impl(h, -1, 0, 0.0, h_out, stream);
}

extern "C" int32_t sb_router_feedback_route(sb_router_t* h, int32_t arm, int64_t rows, double reward, int32_t* h_out,
                                            uintptr_t stream) {
  SB_RANGE("sb_router_feedback_route");
  SB_REQUIRE(arm >= 0, SB_ERR_INVALID, "sb_router_feedback_route: routing %d outside 0..%d", arm, h ? h->n - 1 : 0);
  return router_route_impl(h, arm, rows, reward, h_out, stream);
}

extern "C" int32_t sb_router_last_values(sb_router_t* h, double* h_values, uintptr_t stream) {
  SB_RANGE("sb_router_last_values");
  SB_REQUIRE(h && h_values, SB_ERR_INVALID, "sb_router_last_values: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  SB_CUDA(cudaMemcpyAsync(h_values, h->d_state->last_values, sizeof(double) * h->n, cudaMemcpyDeviceToHost, st));
  SB_CUDA(cudaStreamSynchronize(st));
  return SB_OK;
}
