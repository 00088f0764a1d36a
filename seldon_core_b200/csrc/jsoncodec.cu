// REST JSON payload decode: the text of "ndarray": [[...], ...] -> fp32 device matrix.
//
// Replaces python/seldon_core/utils.py:499-505 (extract_request_parts_json: json.loads, then
// np.array(datadef["ndarray"])) for numeric 2-D (and 1-D) arrays - SURVEY 8(f) row 1, the payload form the engine
// sends over REST.  CPython's json module converts every number with float(), i.e. a correctly rounded
// decimal -> binary64 conversion; the values are then narrowed to fp32 exactly like the Tensor path.
//
// Three kernels over chunks of 4096 bytes (one CTA each, staged in shared memory with coalesced 16-byte loads):
//   json_count  : a token starts where a number character follows a non-number character; counts the tokens, the
//                 brackets and the characters that cannot occur in a numeric array, per chunk
//   json_scan   : exclusive scan of the per-chunk token counts (one CTA)
//   json_parse  : every token start knows its global index k = chunk prefix + rank inside the chunk, hence its
//                 (row, column) = (k / cols, k % cols); the thread that owns the start converts the number and stores
//                 it.  Rectangularity: the token count in front of every inner '[' must be a multiple of cols.
// Conversion: up to 19 significant digits into a 64-bit integer w and a decimal exponent q, then the Eisel-Lemire
// algorithm (128-bit product with a tabulated power of five, csrc/pow5_table.inc from tools/gen_pow5_table.py), which
// is correctly rounded for every (w, q); longer inputs are converted at w and w + 1 and accepted when both agree,
// otherwise the call reports "unsupported number" (no silent approximation, no host fallback).
// Algorithmic traffic: the text once + 4 B per value.
#include "common.cuh"

namespace sb {
namespace {

__device__ const uint64_t kPow5[] = {
#include "pow5_table.inc"
};
constexpr int kSmallestQ = -342, kLargestQ = 308;

constexpr int kJsonChunk = 4096;   // bytes per CTA
constexpr int kJsonLook = 64;      // bytes a number may extend past its chunk (longer tokens are rejected)
constexpr int kJsonPer = kJsonChunk / 256;  // bytes per thread

enum : int32_t { JSON_OK = 0, JSON_MALFORMED = 1, JSON_UNSUPPORTED = 2 };

// characters a numeric token is made of: digits, sign, point, exponent, and the letters of NaN / Infinity
// (membership tests are one shift of a 128-bit set held in four immediates)
constexpr unsigned long long set_lo(const char* s) {
  unsigned long long m = 0;
  for (int i = 0; s[i]; ++i) if ((unsigned char)s[i] < 64) m |= 1ull << s[i];
  return m;
}
constexpr unsigned long long set_hi(const char* s) {
  unsigned long long m = 0;
  for (int i = 0; s[i]; ++i) if ((unsigned char)s[i] >= 64 && (unsigned char)s[i] < 128) m |= 1ull << (s[i] - 64);
  return m;
}
__device__ __forceinline__ bool in_set(uint8_t c, unsigned long long lo, unsigned long long hi) {
  const unsigned long long m = c < 64 ? lo : hi;
  return c < 128 && ((m >> (c & 63)) & 1ull);
}
#define SB_JSON_TOKEN "0123456789-+.eENaInfity"
#define SB_JSON_OTHER "[], \n\r\t"
__device__ __forceinline__ bool token_char(uint8_t c) { return in_set(c, set_lo(SB_JSON_TOKEN), set_hi(SB_JSON_TOKEN)); }
__device__ __forceinline__ bool space_char(uint8_t c) { return c == ' ' || c == '\n' || c == '\r' || c == '\t'; }
// anything else that may appear between the numbers of an array
__device__ __forceinline__ bool struct_char(uint8_t c) { return c == '[' || c == ']' || c == ','; }
__device__ __forceinline__ bool other_char(uint8_t c) { return in_set(c, set_lo(SB_JSON_OTHER), set_hi(SB_JSON_OTHER)); }

// Eisel-Lemire: the binary64 nearest to w * 10^q (w != 0), as raw bits without the sign
__device__ uint64_t decimal_to_double_bits(int64_t q, uint64_t w) {
  if (q < kSmallestQ) return 0ull;
  if (q > kLargestQ) return 0x7ffull << 52;
  const int lz = __clzll((long long)w);
  w <<= lz;
  const uint64_t hi5 = kPow5[2 * (q - kSmallestQ)], lo5 = kPow5[2 * (q - kSmallestQ) + 1];
  uint64_t phi = __umul64hi(w, hi5), plo = w * hi5;
  const uint64_t mask = 0xffffffffffffffffull >> 55;  // 52 + 3 bits of precision wanted
  if ((phi & mask) == mask) {  // the truncated product is too close to a boundary: use the low word of the power too
    const uint64_t shi = __umul64hi(w, lo5);
    plo += shi;
    if (shi > plo) ++phi;
  }
  const int upper = (int)(phi >> 63);
  uint64_t mant = phi >> (upper + 64 - 52 - 3);
  int power2 = (int)((((152170ll + 65536ll) * q) >> 16) + 63) + upper - lz + 1023;
  if (power2 <= 0) {  // subnormal
    if (-power2 + 1 >= 64) return 0ull;
    mant >>= (-power2 + 1);
    mant += mant & 1;
    mant >>= 1;
    return mant;  // the implicit bit, if the rounding produced it, lands in the exponent field: 2^-1022
  }
  // exactly half way: round to even (only possible for a small range of q)
  if (plo <= 1 && q >= -4 && q <= 23 && (mant & 3) == 1 && (mant << (upper + 64 - 52 - 3)) == phi) mant &= ~1ull;
  mant += mant & 1;
  mant >>= 1;
  if (mant >= (2ull << 52)) {
    mant = 1ull << 52;
    ++power2;
  }
  mant &= ~(1ull << 52);
  if (power2 >= 0x7ff) return 0x7ffull << 52;
  return ((uint64_t)power2 << 52) | mant;
}

// converts the token that starts at s[0]; `avail` bytes are readable.  Returns a status and the value.
__device__ int32_t parse_number(const uint8_t* s, int avail, double* out) {
  int p = 0;
  bool neg = false;
  if (s[p] == '-') { neg = true; ++p; }
  if (p + 2 < avail && s[p] == 'N' && s[p + 1] == 'a' && s[p + 2] == 'N' && !neg) {
    p += 3;
    *out = __longlong_as_double(0x7ff8000000000000ll);
  } else if (p + 7 < avail && s[p] == 'I' && s[p + 1] == 'n' && s[p + 2] == 'f' && s[p + 3] == 'i' && s[p + 4] == 'n' &&
             s[p + 5] == 'i' && s[p + 6] == 't' && s[p + 7] == 'y') {
    p += 8;
    *out = __longlong_as_double((long long)((0x7ffull << 52) | (neg ? 1ull << 63 : 0ull)));
  } else {
    uint64_t w = 0;
    int nd = 0, ndig = 0;  // digits kept in w, digits seen
    long long exp10 = 0;
    bool trunc = false;
    for (; p < avail && s[p] >= '0' && s[p] <= '9'; ++p, ++ndig) {
      const uint32_t dgt = s[p] - '0';
      if (nd < 19) { w = w * 10 + dgt; nd += (w != 0); }
      else { ++exp10; trunc |= dgt != 0; }
    }
    if (ndig == 0) return JSON_MALFORMED;
    if (ndig > 1 && s[p - ndig] == '0') return JSON_MALFORMED;  // JSON has no leading zeros ("01" is not a number)
    if (p < avail && s[p] == '.') {
      ++p;
      int nfrac = 0;
      for (; p < avail && s[p] >= '0' && s[p] <= '9'; ++p, ++nfrac) {
        const uint32_t dgt = s[p] - '0';
        if (nd < 19) { w = w * 10 + dgt; nd += (w != 0); --exp10; }
        else trunc |= dgt != 0;
      }
      if (nfrac == 0) return JSON_MALFORMED;
    }
    if (p < avail && (s[p] == 'e' || s[p] == 'E')) {
      ++p;
      bool eneg = false;
      if (p < avail && (s[p] == '+' || s[p] == '-')) { eneg = s[p] == '-'; ++p; }
      long long e = 0;
      int ne = 0;
      for (; p < avail && s[p] >= '0' && s[p] <= '9'; ++p, ++ne)
        if (e < 100000) e = e * 10 + (s[p] - '0');
      if (ne == 0) return JSON_MALFORMED;
      exp10 += eneg ? -e : e;
    }
    uint64_t bits = 0;
    if (w != 0) {
      bits = decimal_to_double_bits(exp10, w);
      if (trunc && decimal_to_double_bits(exp10, w + 1) != bits) return JSON_UNSUPPORTED;
    }
    *out = __longlong_as_double((long long)(bits | (neg ? 1ull << 63 : 0ull)));
  }
  if (p >= avail) return JSON_MALFORMED;  // token longer than the look-ahead window
  const uint8_t t = s[p];
  if (!(struct_char(t) || space_char(t) || t == '}' || t == 0)) return JSON_MALFORMED;
  return JSON_OK;
}

// character classes through a 256-byte table in shared memory: one load per text byte instead of a dozen compares
constexpr uint8_t kClsToken = 1, kClsOther = 2, kClsOpen = 4, kClsClose = 8, kClsComma = 16, kClsSpace = 32;
__device__ __forceinline__ void fill_class_table(uint8_t* lut) {
  for (int c = threadIdx.x; c < 256; c += blockDim.x)
    lut[c] = (uint8_t)((token_char((uint8_t)c) ? kClsToken : 0) | (other_char((uint8_t)c) ? kClsOther : 0) |
                       (c == '[' ? kClsOpen : 0) | (c == ']' ? kClsClose : 0) | (c == ',' ? kClsComma : 0) |
                       (space_char((uint8_t)c) ? kClsSpace : 0));
}
// classes of this thread's kJsonPer (= 16) bytes and of the byte before them
__device__ __forceinline__ void classify16(const uint8_t* t, int i0, const uint8_t* lut, uint8_t (&cls)[kJsonPer],
                                           uint8_t& cls_before) {
  const uint4 w = *reinterpret_cast<const uint4*>(t + i0);
  const uint32_t ww[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
  for (int k = 0; k < kJsonPer; ++k) cls[k] = lut[(ww[k >> 2] >> (8 * (k & 3))) & 0xffu];
  cls_before = lut[t[i0 - 1]];
}

// stage chunk `ch` (+ one byte before, + look-ahead after) into shared memory; bytes outside the text read as 0
__device__ __forceinline__ void stage_chunk(const uint8_t* __restrict__ text, uint64_t nbytes, int64_t ch, uint8_t* sm) {
  // sm[16 + i] = text[ch * kJsonChunk + i] for i in [-1, kJsonChunk + kJsonLook)
  const int64_t base = ch * (int64_t)kJsonChunk - 16;
  const bool aligned = (((uintptr_t)text) & 15) == 0;
  constexpr int NV = (16 + kJsonChunk + kJsonLook) / 16;
  for (int v = threadIdx.x; v < NV; v += blockDim.x) {
    const int64_t o = base + 16 * (int64_t)v;
    uint4 w = make_uint4(0u, 0u, 0u, 0u);
    if (aligned && o >= 0 && o + 16 <= (int64_t)nbytes) {
      w = __ldg(reinterpret_cast<const uint4*>(text + o));
    } else {
      uint8_t* b = reinterpret_cast<uint8_t*>(&w);
      for (int k = 0; k < 16; ++k) b[k] = (o + k >= 0 && o + k < (int64_t)nbytes) ? text[o + k] : (uint8_t)0;
    }
    reinterpret_cast<uint4*>(sm)[v] = w;
  }
}

// per chunk: tokens | '[' << 16 | ']' << 32 | other characters << 48 (each at most 4096)
__global__ void __launch_bounds__(256)
json_count(const uint8_t* __restrict__ text, uint64_t nbytes, int64_t* __restrict__ chunk_info) {
  __shared__ __align__(16) uint8_t sm[16 + kJsonChunk + kJsonLook];
  __shared__ long long warp_info[8];
  __shared__ uint8_t lut[256];
  const int64_t ch = blockIdx.x;
  fill_class_table(lut);
  stage_chunk(text, nbytes, ch, sm);
  __syncthreads();
  const uint8_t* t = sm + 16;
  const int64_t lim = min((int64_t)kJsonChunk, (int64_t)nbytes - ch * (int64_t)kJsonChunk);
  int ntok = 0, nopen = 0, nclose = 0, nbad = 0;
  const int i0 = threadIdx.x * kJsonPer;
  uint8_t cls[kJsonPer], prev;
  classify16(t, i0, lut, cls, prev);
#pragma unroll
  for (int k = 0; k < kJsonPer; ++k) {
    const bool live = i0 + k < lim;
    const uint8_t c = cls[k];
    ntok += live && (c & kClsToken) && !(prev & kClsToken);
    nopen += live && (c & kClsOpen);
    nclose += live && (c & kClsClose);
    nbad += live && !(c & (kClsToken | kClsOther));
    prev = c;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    ntok += __shfl_xor_sync(0xffffffffu, ntok, o);
    nopen += __shfl_xor_sync(0xffffffffu, nopen, o);
    nclose += __shfl_xor_sync(0xffffffffu, nclose, o);
    nbad += __shfl_xor_sync(0xffffffffu, nbad, o);
  }
  if ((threadIdx.x & 31) == 0)
    warp_info[threadIdx.x >> 5] = (long long)ntok | ((long long)nopen << 16) | ((long long)nclose << 32) | ((long long)nbad << 48);
  __syncthreads();
  if (threadIdx.x == 0) {
    long long s = 0;
    for (int w = 0; w < 8; ++w) s += warp_info[w];  // the fields cannot carry into each other (each sum <= 4096)
    chunk_info[ch] = s;
  }
}

// one CTA of 1024 threads: totals of the four counts, and the chunk records replaced by the exclusive prefix of the
// token counts
__global__ void __launch_bounds__(1024)
json_scan(int64_t* __restrict__ chunk_info, int64_t nchunks, int64_t* __restrict__ counts4) {
  __shared__ long long part[1024][4];
  const int64_t per = (nchunks + 1023) / 1024;
  const int64_t a = threadIdx.x * per, b = min(nchunks, a + per);
  long long s[4] = {0, 0, 0, 0};
  for (int64_t i = a; i < b; ++i) {
    const long long v = chunk_info[i];
    s[0] += v & 0xffff; s[1] += (v >> 16) & 0xffff; s[2] += (v >> 32) & 0xffff; s[3] += (v >> 48) & 0xffff;
  }
#pragma unroll
  for (int k = 0; k < 4; ++k) part[threadIdx.x][k] = s[k];
  __syncthreads();
  if (threadIdx.x < 4) {
    long long run = 0;
    for (int i = 0; i < 1024; ++i) { const long long v = part[i][threadIdx.x]; part[i][threadIdx.x] = run; run += v; }
    counts4[threadIdx.x] = run;
  }
  __syncthreads();
  long long run = part[threadIdx.x][0];
  for (int64_t i = a; i < b; ++i) { const long long v = chunk_info[i] & 0xffff; chunk_info[i] = run; run += v; }
}

__global__ void __launch_bounds__(256)
json_parse(const uint8_t* __restrict__ text, uint64_t nbytes, const int64_t* __restrict__ chunk_prefix, int64_t rows,
           int64_t cols, float* __restrict__ out, int64_t ld, int32_t* __restrict__ status) {
  __shared__ __align__(16) uint8_t sm[16 + kJsonChunk + kJsonLook];
  __shared__ int warp_tok[8];
  __shared__ uint8_t lut[256];
  const int64_t ch = blockIdx.x;
  fill_class_table(lut);
  stage_chunk(text, nbytes, ch, sm);
  __syncthreads();
  const uint8_t* t = sm + 16;
  const int64_t lim = min((int64_t)kJsonChunk, (int64_t)nbytes - ch * (int64_t)kJsonChunk);
  const int i0 = threadIdx.x * kJsonPer;
  // token starts and inner '[' among this thread's bytes, as bit masks; `gram` marks every character whose left
  // neighbour (white space skipped) the array grammar constrains: number starts, ',', '[' and ']'
  uint32_t starts = 0, opens = 0, gram = 0;
  {
    uint8_t cls[kJsonPer], prev;
    classify16(t, i0, lut, cls, prev);
#pragma unroll
    for (int k = 0; k < kJsonPer; ++k) {
      const bool live = i0 + k < lim;
      if (live && (cls[k] & kClsToken) && !(prev & kClsToken)) { starts |= 1u << k; gram |= 1u << k; }
      if (live && (cls[k] & kClsOpen) && ch * (int64_t)kJsonChunk + i0 + k != 0) opens |= 1u << k;
      if (live && (cls[k] & (kClsOpen | kClsClose | kClsComma))) gram |= 1u << k;
      prev = cls[k];
    }
  }
  // exclusive scan of the per-thread counts over the CTA
  const int mine = __popc(starts);
  int incl = mine;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int v = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += v;
  }
  if (lane == 31) warp_tok[warp] = incl;
  __syncthreads();
  int before = incl - mine;
  for (int w = 0; w < warp; ++w) before += warp_tok[w];
  int64_t k = chunk_prefix[ch] + before;  // global index of this thread's first token
  int32_t bad = JSON_OK;
  // The array grammar, as a condition on the nearest non-blank character to the left of every number start and of
  // every structural character (what json.loads enforces, utils.py:499-505 -> json.loads):
  //   number : '[' or ','          ','  : the end of a number or ']'
  //   '['    : '[' or ',' (or the first character)      ']' : the end of a number, ']' or '[' (an empty list)
  // Together with the bracket counts checked by the caller this rejects "[[1 2]]", "[[1,,2]]", "[1,2,]", "[,1]", ...
  while (gram) {
    const int j = __ffs((int)gram) - 1;
    gram &= gram - 1;
    const int i = i0 + j;
    const int64_t gpos = ch * (int64_t)kJsonChunk + i;
    if (gpos == 0) continue;  // the opening bracket of the whole array
    int q = i - 1;
    uint8_t pc = 0;
    bool found = false;
    for (; q >= -16; --q) {
      pc = t[q];
      if (!(lut[pc] & kClsSpace)) { found = true; break; }
    }
    if (!found) {  // a run of more than 16 blanks reaches back past the staged bytes: read the text itself
      for (int64_t g = ch * (int64_t)kJsonChunk - 17; g >= 0; --g) {
        pc = text[g];
        if (!(lut[pc] & kClsSpace)) { found = true; break; }
      }
    }
    const uint8_t me = lut[t[i]], pl = found ? lut[pc] : 0;
    const bool p_open = pl & kClsOpen, p_close = pl & kClsClose, p_comma = pl & kClsComma, p_num = pl & kClsToken;
    bool ok;
    if (me & kClsToken) ok = p_open || p_comma;
    else if (me & kClsComma) ok = p_num || p_close;
    else if (me & kClsOpen) ok = p_open || p_comma;
    else ok = p_num || p_close || p_open;
    if (!found || !ok) bad = JSON_MALFORMED;
  }
  // an inner row opens at every '[' but the first: everything before it must be whole rows
  while (opens) {
    const int j = __ffs((int)opens) - 1;
    opens &= opens - 1;
    const int64_t kk = k + __popc(starts & ((1u << j) - 1u));
    if (cols == 0 || kk % cols != 0 || kk / cols >= rows) bad = JSON_MALFORMED;
  }
  // the lanes of a warp convert their n-th token together (no serialisation over the byte position)
  while (starts) {
    const int j = __ffs((int)starts) - 1;
    starts &= starts - 1;
    const int i = i0 + j;
    double v = 0.0;
    const int32_t rc = parse_number(t + i, kJsonChunk + kJsonLook - i, &v);
    bad = max(bad, rc);
    if (k < rows * cols) out[(k / cols) * ld + (k % cols)] = (float)v;
    else bad = max(bad, (int32_t)JSON_MALFORMED);
    ++k;
  }
  if (bad != JSON_OK) atomicMax(status, bad);
}

}  // namespace
}  // namespace sb

using namespace sb;

extern "C" uint64_t sb_json_workspace_bytes(uint64_t nbytes) {
  const uint64_t nchunks = (nbytes + kJsonChunk - 1) / kJsonChunk;
  return (nchunks + 1) * sizeof(int64_t) + 4 * sizeof(unsigned long long);
}

extern "C" int32_t sb_json_ndarray_count(const uint8_t* d_text, uint64_t nbytes, void* d_ws, uint64_t ws_bytes,
                                         int64_t* d_counts4, uintptr_t stream) {
  SB_RANGE("sb_json_ndarray_count");
  SB_REQUIRE(d_text && nbytes >= 2 && d_ws && d_counts4, SB_ERR_INVALID, "sb_json_ndarray_count: bad arguments");
  SB_REQUIRE(((uintptr_t)d_ws & 7) == 0 && ws_bytes >= sb_json_workspace_bytes(nbytes), SB_ERR_WORKSPACE,
             "sb_json_ndarray_count: workspace too small or misaligned");
  const int64_t nchunks = (int64_t)((nbytes + kJsonChunk - 1) / kJsonChunk);
  SB_REQUIRE(nchunks <= 0x7fffffff, SB_ERR_INVALID, "sb_json_ndarray_count: text too long for one call");
  cudaStream_t st = (cudaStream_t)stream;
  int64_t* chunk_tokens = (int64_t*)d_ws;
  json_count<<<(unsigned)nchunks, 256, 0, st>>>(d_text, nbytes, chunk_tokens);
  SB_LAUNCH_CHECK();
  json_scan<<<1, 1024, 0, st>>>(chunk_tokens, nchunks, d_counts4);
  SB_LAUNCH_CHECK();
  return SB_OK;
}

extern "C" int32_t sb_unpack_json_ndarray(const uint8_t* d_text, uint64_t nbytes, int64_t rows, int64_t cols,
                                          const void* d_ws, float* d_out, int64_t ld, int32_t* d_status,
                                          uintptr_t stream) {
  SB_RANGE("sb_unpack_json_ndarray");
  SB_REQUIRE(d_text && nbytes >= 2 && d_ws && d_status, SB_ERR_INVALID, "sb_unpack_json_ndarray: bad arguments");
  SB_REQUIRE(rows >= 0 && cols >= 0 && ld >= cols, SB_ERR_INVALID, "sb_unpack_json_ndarray: bad shape");
  SB_REQUIRE(rows * cols == 0 || d_out, SB_ERR_INVALID, "sb_unpack_json_ndarray: null output");
  cudaStream_t st = (cudaStream_t)stream;
  SB_CUDA(cudaMemsetAsync(d_status, 0, sizeof(int32_t), st));
  const int64_t nchunks = (int64_t)((nbytes + kJsonChunk - 1) / kJsonChunk);
  json_parse<<<(unsigned)nchunks, 256, 0, st>>>(d_text, nbytes, (const int64_t*)d_ws, rows, cols, d_out, ld, d_status);
  SB_LAUNCH_CHECK();
  return SB_OK;
}
