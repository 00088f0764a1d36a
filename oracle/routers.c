/* C twin of oracle/routers.py - TEST INFRASTRUCTURE ONLY (never linked into libsb200.so).
 *
 * Restates EpsilonGreedy.py:35-136 and ThompsonSampling.py:35-115 of the reference together with the two
 * MT19937 streams they draw from (CPython `random`, NumPy legacy RandomState); see routers.py for the
 * third-party sources restated.  Scalar, single thread; used to check the 10 M-call configuration.
 */
#include <math.h>
#include <stdint.h>
#include <string.h>

#define MT_N 624
#define MT_M 397

typedef struct {
  uint32_t mt[MT_N];
  int pos;
  int has_gauss;
  double gauss;
} mt_t;

static void mt_init_genrand(mt_t* g, uint32_t s) {
  g->mt[0] = s;
  for (int i = 1; i < MT_N; ++i) g->mt[i] = 1812433253u * (g->mt[i - 1] ^ (g->mt[i - 1] >> 30)) + (uint32_t)i;
  g->pos = MT_N;
  g->has_gauss = 0;
  g->gauss = 0.0;
}

static void mt_init_by_array(mt_t* g, const uint32_t* key, int len) {
  mt_init_genrand(g, 19650218u);
  uint32_t* mt = g->mt;
  int i = 1, j = 0;
  int k = MT_N > len ? MT_N : len;
  for (; k; --k) {
    mt[i] = (mt[i] ^ ((mt[i - 1] ^ (mt[i - 1] >> 30)) * 1664525u)) + key[j] + (uint32_t)j;
    ++i; ++j;
    if (i >= MT_N) { mt[0] = mt[MT_N - 1]; i = 1; }
    if (j >= len) j = 0;
  }
  for (k = MT_N - 1; k; --k) {
    mt[i] = (mt[i] ^ ((mt[i - 1] ^ (mt[i - 1] >> 30)) * 1566083941u)) - (uint32_t)i;
    ++i;
    if (i >= MT_N) { mt[0] = mt[MT_N - 1]; i = 1; }
  }
  mt[0] = 0x80000000u;
  g->pos = MT_N;
}

static uint32_t mt_u32(mt_t* g) {
  if (g->pos >= MT_N) {
    uint32_t* mt = g->mt;
    for (int k = 0; k < MT_N; ++k) {
      uint32_t y = (mt[k] & 0x80000000u) | (mt[(k + 1) % MT_N] & 0x7fffffffu);
      mt[k] = mt[(k + MT_M) % MT_N] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
    }
    g->pos = 0;
  }
  uint32_t y = g->mt[g->pos++];
  y ^= y >> 11;
  y ^= (y << 7) & 0x9d2c5680u;
  y ^= (y << 15) & 0xefc60000u;
  y ^= y >> 18;
  return y;
}

static double mt_double(mt_t* g) {
  uint32_t a = mt_u32(g) >> 5, b = mt_u32(g) >> 6;
  return (a * 67108864.0 + b) / 9007199254740992.0;
}

static uint32_t py_randbelow(mt_t* g, uint32_t n) {
  int k = 32 - __builtin_clz(n);
  uint32_t r = mt_u32(g) >> (32 - k);
  while (r >= n) r = mt_u32(g) >> (32 - k);
  return r;
}

static uint32_t np_choice_index(mt_t* g, uint32_t m) {
  uint32_t rng = m - 1;
  if (rng == 0) return 0;
  uint32_t mask = rng;
  mask |= mask >> 1; mask |= mask >> 2; mask |= mask >> 4; mask |= mask >> 8; mask |= mask >> 16;
  for (;;) {
    uint32_t v = mt_u32(g) & mask;
    if (v <= rng) return v;
  }
}

static double np_gauss(mt_t* g) {
  if (g->has_gauss) {
    double t = g->gauss;
    g->has_gauss = 0;
    g->gauss = 0.0;
    return t;
  }
  double x1, x2, r2;
  do {
    x1 = 2.0 * mt_double(g) - 1.0;
    x2 = 2.0 * mt_double(g) - 1.0;
    r2 = x1 * x1 + x2 * x2;
  } while (r2 >= 1.0 || r2 == 0.0);
  double f = sqrt(-2.0 * log(r2) / r2);
  g->gauss = f * x1;
  g->has_gauss = 1;
  return f * x2;
}

static double np_std_exponential(mt_t* g) { return -log(1.0 - mt_double(g)); }

static double np_std_gamma(mt_t* g, double shape) {
  if (shape == 1.0) return np_std_exponential(g);
  if (shape == 0.0) return 0.0;
  if (shape < 1.0) {
    for (;;) {
      double u = mt_double(g), v = np_std_exponential(g);
      if (u <= 1.0 - shape) {
        double x = pow(u, 1.0 / shape);
        if (x <= v) return x;
      } else {
        double y = -log((1 - u) / shape);
        double x = pow(1.0 - shape + shape * y, 1.0 / shape);
        if (x <= v + y) return x;
      }
    }
  }
  double b = shape - 1.0 / 3.0, c = 1.0 / sqrt(9 * b);
  for (;;) {
    double x, v;
    do {
      x = np_gauss(g);
      v = 1.0 + c * x;
    } while (v <= 0.0);
    v = v * v * v;
    double u = mt_double(g);
    if (u < 1.0 - 0.0331 * (x * x) * (x * x)) return b * v;
    if (log(u) < 0.5 * x * x + b * (1.0 - v + log(v))) return b * v;
  }
}

static double np_beta(mt_t* g, double a, double b) {
  if (a <= 1.0 && b <= 1.0) {
    for (;;) {
      double u = mt_double(g), v = mt_double(g);
      double x = pow(u, 1.0 / a), y = pow(v, 1.0 / b);
      double xpy = x + y;
      if (xpy <= 1.0 && u + v > 0.0) {
        if (xpy > 0) return x / xpy;
        double lx = log(u) / a, ly = log(v) / b;
        double lm = lx > ly ? lx : ly;
        lx -= lm; ly -= lm;
        return exp(lx - log(exp(lx) + exp(ly)));
      }
    }
  }
  double ga = np_std_gamma(g, a), gb = np_std_gamma(g, b);
  return ga / (ga + gb);
}

#define MAX_ARMS 1024

/* kind 0 = epsilon-greedy, 1 = Thompson.  T interleaved route_t / send_feedback_t calls.
 * reward_mode 0: reward[t]; 1: reward[t*n+arm]; 2: reward[t] is u_t, reward = u_t < arm_p[arm] ? 1 : 0.
 * state out: success[n], tries[n], best_branch.  Returns 0, or -1 on bad arguments. */
int sbo_router_run(int kind, int n, double epsilon, int64_t seed, int best_branch, int64_t T, const int64_t* rows,
                   const float* reward, const float* arm_p, int reward_mode, int32_t* decisions, int64_t* state) {
  if (n <= 0 || n > MAX_ARMS || seed <= 0) return -1;
  static mt_t py, np;
  int64_t succ[MAX_ARMS], tries[MAX_ARMS];
  double values[MAX_ARMS];
  memset(succ, 0, sizeof succ);
  memset(tries, 0, sizeof tries);
  memset(values, 0, sizeof values);
  {
    uint32_t key[2];
    int len = 0;
    uint64_t a = (uint64_t)seed;
    while (a) { key[len++] = (uint32_t)a; a >>= 32; }
    mt_init_by_array(&py, key, len);
    mt_init_genrand(&np, (uint32_t)seed);
  }
  int best = 0;
  if (kind == 0) best = best_branch ? best_branch : (int)py_randbelow(&py, (uint32_t)n);
  for (int64_t t = 0; t < T; ++t) {
    int arm;
    if (kind == 0) {
      double x = mt_double(&py);
      if (x > epsilon) arm = best;
      else {
        int r = (int)py_randbelow(&py, (uint32_t)(n - 1));
        arm = r < best ? r : r + 1;
      }
    } else {
      double top = 0.0;
      arm = 0;
      for (int j = 0; j < n; ++j) {
        double a = (double)(1 + succ[j]), b = (double)(1 + tries[j] - succ[j]);
        double v = np_beta(&np, a, b);
        if (j == 0 || v > top) { top = v; arm = j; }
      }
    }
    decisions[t] = arm;
    int64_t nrows = rows ? rows[t] : 1;
    double r;
    if (reward_mode == 0) r = (double)reward[t];
    else if (reward_mode == 1) r = (double)reward[t * n + arm];
    else r = reward[t] < arm_p[arm] ? 1.0 : 0.0;
    int64_t ns = (int64_t)(r * (double)nrows);
    succ[arm] += ns;
    tries[arm] += nrows;
    if (kind == 0) {
      values[arm] = (double)succ[arm] / (double)tries[arm];
      double top = values[0];
      for (int j = 1; j < n; ++j) if (values[j] > top) top = values[j];
      int m = 0;
      for (int j = 0; j < n; ++j) if (values[j] == top) ++m;
      int pick = (int)np_choice_index(&np, (uint32_t)m);
      for (int j = 0; j < n; ++j)
        if (values[j] == top && pick-- == 0) { best = j; break; }
    }
  }
  for (int j = 0; j < n; ++j) { state[j] = succ[j]; state[n + j] = tries[j]; }
  state[2 * n] = best;
  return 0;
}
