// digest_device.cuh — sm_100a device code for the byte-integrity kernels.
//
// One thread = one message (Merkle–Damgård chains are sequential), one warp = 32 messages.
// Each lane's message is staged HBM -> shared memory with 1-D bulk async copies
// (cp.async.bulk, SASS UBLKCP) that complete on a per-warp mbarrier, double/triple buffered, and
// read back as LDS.128.  The compression functions are written so ptxas emits LOP3 for every
// 3-input boolean and SHF for every rotate.
//
// What it replaces (there is no reference source for the arithmetic — it is OpenSSL behind Node
// crypto): sha1 per piece ([dep] simple-sha1@2.1.2 via webtorrent, call site lib/download.js:64-66),
// md5 / sha256 per part ([dep] minio@7.0.12 getHashSummer, call site lib/upload.js:45,55).
// Algorithms: FIPS 180-4 §6.1 (SHA-1), §6.2 (SHA-256), RFC 1321 (MD5).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace th {

// ------------------------------------------------------------------------------------ primitives
template <int LUT>
__device__ __forceinline__ uint32_t lop3(uint32_t a, uint32_t b, uint32_t c) {
  uint32_t r;
  asm("lop3.b32 %0, %1, %2, %3, %4;" : "=r"(r) : "r"(a), "r"(b), "r"(c), "n"(LUT));
  return r;
}
__device__ __forceinline__ uint32_t rotl(uint32_t x, int n) { return __funnelshift_l(x, x, n); }
__device__ __forceinline__ uint32_t rotr(uint32_t x, int n) { return __funnelshift_r(x, x, n); }
__device__ __forceinline__ uint32_t bswap(uint32_t x) { return __byte_perm(x, 0, 0x0123); }

// a + b issued as IMAD on the FMA pipe instead of IADD3 on the (saturated) ALU pipe.  `one` is the
// kernel argument 1: a literal would be folded back into an add.  The alu pipe issues one warp
// instruction every 2 cycles per sub-partition and so does the fma pipe, so adds moved here run
// beside the LOP3/SHF work for free (B300_MICROARCH.md "Pipe rates").
__device__ __forceinline__ uint32_t add_fma(uint32_t a, uint32_t b, uint32_t one) {
  uint32_t r;
  asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(one), "r"(b));
  return r;
}

constexpr int LUT_CH = 0xCA;   // (a & b) | (~a & c)
constexpr int LUT_MAJ = 0xE8;  // majority
constexpr int LUT_XOR3 = 0x96; // a ^ b ^ c
constexpr int LUT_MD5I = 0x39; // b ^ (a | ~c)   with (a,b,c) = (b,c,d) of RFC 1321's I(b,c,d) = c ^ (b | ~d)

// ---------------------------------------------------------------------------------------- SHA-1
// SLICED_BOUND: the thread bound given to ptxas for the persistent kernel (it runs 384 threads).  Only the
// schedule ptxas picks changes with it; measured on B200: SHA-256 919 GB/s at 512, 968 at 704..768, 884 at 896.
struct Sha1 {
  static constexpr int NS = 5, DLEN = 20, SLICED_BOUND = 512;
  static constexpr bool BE = true;
  __device__ static __forceinline__ void init(uint32_t* h) {
    h[0] = 0x67452301u; h[1] = 0xefcdab89u; h[2] = 0x98badcfeu; h[3] = 0x10325476u; h[4] = 0xc3d2e1f0u;
  }
  // w: 16 big-endian-decoded words; used as the rolling schedule window (destroyed).
  __device__ static __forceinline__ void compress(uint32_t* h, uint32_t* w, uint32_t one) {
    uint32_t a = h[0], b = h[1], c = h[2], d = h[3], e = h[4];
#pragma unroll
    for (int t = 0; t < 80; t++) {
      if (t >= 16)
        w[t & 15] = rotl(lop3<LUT_XOR3>(w[(t - 3) & 15], w[(t - 8) & 15], w[(t - 14) & 15]) ^ w[t & 15], 1);
      uint32_t f, k;
      if (t < 20)      { f = lop3<LUT_CH>(b, c, d);   k = 0x5a827999u; }
      else if (t < 40) { f = lop3<LUT_XOR3>(b, c, d); k = 0x6ed9eba1u; }
      else if (t < 60) { f = lop3<LUT_MAJ>(b, c, d);  k = 0x8f1bbcdcu; }
      else             { f = lop3<LUT_XOR3>(b, c, d); k = 0xca62c1d6u; }
      // x = f + e + K + W[t] is off the critical path (f uses a from two rounds back): three IMADs;
      // the round's only dependent op is rotl(a,5) + x, which ptxas fuses into one LEA.HI
      uint32_t x = add_fma(w[t & 15], k, one);
      x = add_fma(e, x, one);
      x = add_fma(f, x, one);
      uint32_t tmp = rotl(a, 5) + x;
      e = d; d = c; c = rotl(b, 30); b = a; a = tmp;
    }
    h[0] += a; h[1] += b; h[2] += c; h[3] += d; h[4] += e;
  }
};

// ------------------------------------------------------------------------------------------ MD5
struct Md5 {
  static constexpr int NS = 4, DLEN = 16, SLICED_BOUND = 512;
  static constexpr bool BE = false;
  __device__ static __forceinline__ void init(uint32_t* h) {
    h[0] = 0x67452301u; h[1] = 0xefcdab89u; h[2] = 0x98badcfeu; h[3] = 0x10325476u;
  }
  template <int I>
  __device__ static __forceinline__ void step(uint32_t& a, uint32_t b, uint32_t c, uint32_t d,
                                              const uint32_t* m, uint32_t one) {
    constexpr uint32_t T[64] = {
        0xd76aa478, 0xe8c7b756, 0x242070db, 0xc1bdceee, 0xf57c0faf, 0x4787c62a, 0xa8304613, 0xfd469501,
        0x698098d8, 0x8b44f7af, 0xffff5bb1, 0x895cd7be, 0x6b901122, 0xfd987193, 0xa679438e, 0x49b40821,
        0xf61e2562, 0xc040b340, 0x265e5a51, 0xe9b6c7aa, 0xd62f105d, 0x02441453, 0xd8a1e681, 0xe7d3fbc8,
        0x21e1cde6, 0xc33707d6, 0xf4d50d87, 0x455a14ed, 0xa9e3e905, 0xfcefa3f8, 0x676f02d9, 0x8d2a4c8a,
        0xfffa3942, 0x8771f681, 0x6d9d6122, 0xfde5380c, 0xa4beea44, 0x4bdecfa9, 0xf6bb4b60, 0xbebfbc70,
        0x289b7ec6, 0xeaa127fa, 0xd4ef3085, 0x04881d05, 0xd9d4d039, 0xe6db99e5, 0x1fa27cf8, 0xc4ac5665,
        0xf4292244, 0x432aff97, 0xab9423a7, 0xfc93a039, 0x655b59c3, 0x8f0ccc92, 0xffeff47d, 0x85845dd1,
        0x6fa87e4f, 0xfe2ce6e0, 0xa3014314, 0x4e0811a1, 0xf7537e82, 0xbd3af235, 0x2ad7d2bb, 0xeb86d391};
    constexpr int S[16] = {7, 12, 17, 22, 5, 9, 14, 20, 4, 11, 16, 23, 6, 10, 15, 21};
    constexpr int R = I >> 4;
    constexpr int G = R == 0 ? I : R == 1 ? (5 * I + 1) & 15 : R == 2 ? (3 * I + 5) & 15 : (7 * I) & 15;
    uint32_t f = R == 0   ? lop3<LUT_CH>(b, c, d)
                 : R == 1 ? lop3<LUT_CH>(d, b, c)
                 : R == 2 ? lop3<LUT_XOR3>(b, c, d)
                          : lop3<LUT_MD5I>(b, c, d);
    // m[G] + T[I] depends on nothing in the chain: one IMAD on the FMA pipe.  The rest stays on the ALU pipe
    // as IADD3 (a + mk + f) and LEA.HI (rotl + b): 3 ALU + 1 FMA instructions per step.  (All three adds as
    // IMAD measured slower: the step's dependent chain got longer than three warps can hide; moving only the
    // a + mk + f sum to IMADs, LEA.HI kept, also measured slower: 3.9 vs 4.7 TB/s, with 12 or 16 warps per SM.)
    uint32_t akm = a + add_fma(m[G], T[I], one);
    a = b + rotl(akm + f, S[(R << 2) | (I & 3)]);
  }
  template <int I>
  __device__ static __forceinline__ void four(uint32_t& a, uint32_t& b, uint32_t& c, uint32_t& d,
                                              const uint32_t* m, uint32_t one) {
    step<I>(a, b, c, d, m, one);
    step<I + 1>(d, a, b, c, m, one);
    step<I + 2>(c, d, a, b, m, one);
    step<I + 3>(b, c, d, a, m, one);
  }
  // w: 16 little-endian words (not modified).
  __device__ static __forceinline__ void compress(uint32_t* h, const uint32_t* m, uint32_t one) {
    uint32_t a = h[0], b = h[1], c = h[2], d = h[3];
    four<0>(a, b, c, d, m, one);  four<4>(a, b, c, d, m, one);  four<8>(a, b, c, d, m, one);  four<12>(a, b, c, d, m, one);
    four<16>(a, b, c, d, m, one); four<20>(a, b, c, d, m, one); four<24>(a, b, c, d, m, one); four<28>(a, b, c, d, m, one);
    four<32>(a, b, c, d, m, one); four<36>(a, b, c, d, m, one); four<40>(a, b, c, d, m, one); four<44>(a, b, c, d, m, one);
    four<48>(a, b, c, d, m, one); four<52>(a, b, c, d, m, one); four<56>(a, b, c, d, m, one); four<60>(a, b, c, d, m, one);
    h[0] += a; h[1] += b; h[2] += c; h[3] += d;
  }
};

// -------------------------------------------------------------------------------------- SHA-256
struct Sha256 {
  static constexpr int NS = 8, DLEN = 32, SLICED_BOUND = 768;
  static constexpr bool BE = true;
  __device__ static __forceinline__ void init(uint32_t* h) {
    h[0] = 0x6a09e667u; h[1] = 0xbb67ae85u; h[2] = 0x3c6ef372u; h[3] = 0xa54ff53au;
    h[4] = 0x510e527fu; h[5] = 0x9b05688cu; h[6] = 0x1f83d9abu; h[7] = 0x5be0cd19u;
  }
  __device__ static __forceinline__ void compress(uint32_t* h, uint32_t* w, uint32_t one) {
    constexpr uint32_t K[64] = {
        0x428a2f98, 0x71374491, 0xb5c0fbcf, 0xe9b5dba5, 0x3956c25b, 0x59f111f1, 0x923f82a4, 0xab1c5ed5,
        0xd807aa98, 0x12835b01, 0x243185be, 0x550c7dc3, 0x72be5d74, 0x80deb1fe, 0x9bdc06a7, 0xc19bf174,
        0xe49b69c1, 0xefbe4786, 0x0fc19dc6, 0x240ca1cc, 0x2de92c6f, 0x4a7484aa, 0x5cb0a9dc, 0x76f988da,
        0x983e5152, 0xa831c66d, 0xb00327c8, 0xbf597fc7, 0xc6e00bf3, 0xd5a79147, 0x06ca6351, 0x14292967,
        0x27b70a85, 0x2e1b2138, 0x4d2c6dfc, 0x53380d13, 0x650a7354, 0x766a0abb, 0x81c2c92e, 0x92722c85,
        0xa2bfe8a1, 0xa81a664b, 0xc24b8b70, 0xc76c51a3, 0xd192e819, 0xd6990624, 0xf40e3585, 0x106aa070,
        0x19a4c116, 0x1e376c08, 0x2748774c, 0x34b0bcb5, 0x391c0cb3, 0x4ed8aa4a, 0x5b9cca4f, 0x682e6ff3,
        0x748f82ee, 0x78a5636f, 0x84c87814, 0x8cc70208, 0x90befffa, 0xa4506ceb, 0xbef9a3f7, 0xc67178f2};
    uint32_t a = h[0], b = h[1], c = h[2], d = h[3], e = h[4], f = h[5], g = h[6], hh = h[7];
#pragma unroll
    for (int t = 0; t < 64; t++) {
      if (t >= 16) {
        uint32_t x = w[(t - 15) & 15], y = w[(t - 2) & 15];
        uint32_t s0 = lop3<LUT_XOR3>(rotr(x, 7), rotr(x, 18), x >> 3);
        uint32_t s1 = lop3<LUT_XOR3>(rotr(y, 17), rotr(y, 19), y >> 10);
        w[t & 15] = add_fma(add_fma(w[t & 15], s0, one), add_fma(w[(t - 7) & 15], s1, one), one);
      }
      // every add goes to the FMA pipe as IMAD; the ALU pipe keeps only the rotates and booleans
      uint32_t S1 = lop3<LUT_XOR3>(rotr(e, 6), rotr(e, 11), rotr(e, 25));
      uint32_t t1 = add_fma(w[t & 15], K[t], one);  // off the critical path
      t1 = add_fma(hh, t1, one);
      t1 = add_fma(add_fma(S1, lop3<LUT_CH>(e, f, g), one), t1, one);
      uint32_t S0 = lop3<LUT_XOR3>(rotr(a, 2), rotr(a, 13), rotr(a, 22));
      uint32_t t2 = add_fma(S0, lop3<LUT_MAJ>(a, b, c), one);
      hh = g; g = f; f = e; e = add_fma(d, t1, one); d = c; c = b; b = a; a = add_fma(t1, t2, one);
    }
    h[0] += a; h[1] += b; h[2] += c; h[3] += d; h[4] += e; h[5] += f; h[6] += g; h[7] += hh;
  }
};

// ------------------------------------------------------------------- shared memory / TMA helpers
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_init_fence() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "TH_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra TH_DONE;\n"
      "bra TH_WAIT;\n"
      "TH_DONE:\n"
      "}\n" ::"r"(bar),
      "r"(parity)
      : "memory");
}
// 1-D bulk async copy global -> this CTA's shared memory, completing `bytes` on the mbarrier.
// dst, src 16-byte aligned; bytes a non-zero multiple of 16.
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}
// (measured: dropping this fence changes nothing, so it stays)
__device__ __forceinline__ void fence_proxy_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ uint4 lds128(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts8(uint32_t addr, uint32_t v) {
  asm volatile("st.shared.u8 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}

// ------------------------------------------------------------------------------- padding blocks
// Turns the 16 little-endian-loaded words of the block that holds the message tail (r = len % 64
// valid bytes, anything after them garbage) into padding block `k` (0 or 1) of the message:
// 0x80, zeros, 64-bit bit length — little-endian for MD5, big-endian for SHA-1/SHA-256
// (RFC 1321 §3.1-3.2, FIPS 180-4 §5.1.1).  For BE digests the length words are pre-swapped because
// the caller byte-swaps all 16 words afterwards.
template <class A>
__device__ __forceinline__ void pad_block(uint32_t* w, uint32_t k, uint32_t r, uint64_t len) {
  const bool one_block = r < 56;
  if (k == 0) {
    const uint32_t wi = r >> 2, sh = (r & 3) * 8;
#pragma unroll
    for (int i = 0; i < 16; i++) {
      uint32_t keep = (uint32_t)i < wi ? w[i] : 0u;
      uint32_t edge = (w[i] & ((1u << sh) - 1u)) | (0x80u << sh);
      w[i] = (uint32_t)i == wi ? edge : keep;
    }
  } else {
#pragma unroll
    for (int i = 0; i < 16; i++) w[i] = 0u;
  }
  if (k == 1 || one_block) {
    const uint64_t bits = len << 3;
    const uint32_t lo = (uint32_t)bits, hi = (uint32_t)(bits >> 32);
    if (A::BE) { w[14] = bswap(hi); w[15] = bswap(lo); }
    else       { w[14] = lo;        w[15] = hi; }
  }
}

// ------------------------------------------------------------------------------ the stream engine
// Per-warp staging geometry.  CH = bytes staged per lane per stage (multiple of 64), NST = stages.
// Lane slots are CH+16 bytes apart so that the eight 16-byte reads of each LDS.128 quarter-warp
// phase fall in eight different 16-byte bank groups (stride = 16 mod 128 when CH % 128 == 0, an odd
// multiple of 16 otherwise is also conflict-free).
template <int CH, int NST>
struct Staging {
  static_assert(CH % 64 == 0 && CH >= 64 && CH <= 1024, "chunk must be a multiple of 64");
  static_assert(((CH + 16) / 16) % 2 == 1, "lane stride must be an odd multiple of 16 bytes");
  static constexpr int STRIDE = CH + 16;
  static constexpr int STAGE_BYTES = 32 * STRIDE;
  static constexpr int WARP_BYTES = NST * STAGE_BYTES + 8 * NST;  // + mbarriers
};

// Hashes one message per lane.  msg is 16-byte aligned (or anything if len == 0).
// smem: this warp's Staging<CH,NST>::WARP_BYTES region (16-byte aligned).
template <class A, int CH, int NST>
__device__ __forceinline__ void digest_stream(const uint8_t* __restrict__ msg, uint64_t len, uint32_t* h,
                                              unsigned char* smem, int lane, uint32_t one) {
  using G = Staging<CH, NST>;
  constexpr unsigned FULL = 0xffffffffu;
  const uint32_t s_base = smem_u32(smem);
  const uint32_t s_bar = s_base + NST * G::STAGE_BYTES;
  const uint32_t s_lane = s_base + lane * G::STRIDE;

  if (lane == 0) {
#pragma unroll
    for (int s = 0; s < NST; s++) mbar_init(s_bar + 8 * s, 1);
    mbar_init_fence();
  }
  __syncwarp();

  // chunk c covers message bytes [c*CH, min((c+1)*CH, len)); every lane has nch >= 1 chunks and its
  // last chunk is partial (possibly empty) and is followed by the padding block(s).
  const uint32_t nch = (uint32_t)(len / CH) + 1u;
  const uint32_t last_bytes = (uint32_t)(len - (uint64_t)(nch - 1u) * CH);
  const uint32_t max_nch = __reduce_max_sync(FULL, nch);

  auto issue = [&](uint32_t c, int s) {
    uint32_t bytes = c < nch ? (c == nch - 1u ? last_bytes : (uint32_t)CH) : 0u;
    const uint32_t tb = bytes & ~15u;
    const uint32_t total = __reduce_add_sync(FULL, tb);
    if (c < max_nch) {  // warp-uniform
      const uint32_t bar = s_bar + 8 * s;
      const uint32_t dst = s_lane + s * G::STAGE_BYTES;
      const uint8_t* src = msg + (uint64_t)c * CH;
      if (lane == 0) mbar_arrive_expect_tx(bar, total);
      __syncwarp();
      if (tb) bulk_g2s(dst, src, tb, bar);
      // a tail shorter than 16 bytes cannot be bulk-copied: move it by hand (last chunk only)
      for (uint32_t j = tb; j < bytes; j++) sts8(dst + j, __ldg(src + j));
    }
  };

#pragma unroll
  for (int s = 0; s < NST; s++) issue((uint32_t)s, s);

  A::init(h);
  int s = 0;
  uint32_t parity = 0;
  for (uint32_t c = 0; c < max_nch; c++) {
    mbar_wait(s_bar + 8 * s, parity);
    if (c < nch) {
      const bool last = c == nch - 1u;
      const uint32_t bytes = last ? last_bytes : (uint32_t)CH;
      const uint32_t nb = bytes >> 6;                      // whole data blocks in this chunk
      const uint32_t r = bytes & 63u;                      // tail bytes (last chunk only)
      const uint32_t nblk = nb + (last ? (r < 56 ? 1u : 2u) : 0u);
      const uint32_t src = s_lane + s * G::STAGE_BYTES;
      for (uint32_t b = 0; b < nblk; b++) {
        uint32_t w[16];
        // padding block 1 may lie one block past the slot when the tail sits in the slot's last
        // block; it is all zeros + length, so nothing needs to be read for it
        const uint32_t a = src + (b < (uint32_t)(CH / 64) ? b : 0u) * 64u;
#pragma unroll
        for (int q = 0; q < 4; q++) {
          uint4 v = lds128(a + 16 * q);
          w[4 * q] = v.x; w[4 * q + 1] = v.y; w[4 * q + 2] = v.z; w[4 * q + 3] = v.w;
        }
        if (b >= nb) pad_block<A>(w, b - nb, r, len);
        if (A::BE) {
#pragma unroll
          for (int i = 0; i < 16; i++) w[i] = bswap(w[i]);
        }
        A::compress(h, w, one);
      }
    }
    __syncwarp();
    // every value read from this stage has been consumed by compress(); order those generic-proxy
    // reads before the async-proxy refill of the same bytes
    fence_proxy_async_smem();
    issue(c + NST, s);
    if (++s == NST) { s = 0; parity ^= 1u; }
  }
}

// ------------------------------------------------------------------- 2-D tensor-map (TMA) staging
// Equal-length messages a fixed stride apart form a 2-D byte tensor [n rows][len bytes].  One
// cp.async.bulk.tensor.2d (SASS UTMALDG) then moves a {128 B x 32 rows} box — the next 128 bytes of
// all 32 messages of a warp — into a 4 KiB shared-memory tile with the 128-byte swizzle, so lane r
// reads its 16-byte piece q at tile + r*128 + ((q ^ (r & 7)) << 4): the eight lanes of every LDS.128
// quarter-warp phase hit eight different bank groups.  Bytes past `len` and rows past n are
// zero-filled by the TMA unit (no over-read); one elected lane issues, nobody loops.
__device__ __forceinline__ void tma_tile_2d(uint32_t dst, const void* tmap, uint32_t c0, uint32_t c1, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(dst),
      "l"(tmap), "r"(c0), "r"(c1), "r"(bar)
      : "memory");
}
__device__ __forceinline__ void prefetch_tmap(const void* tmap) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(tmap) : "memory");
}

template <int K, int NST>
struct Tiles {
  static_assert(K >= 1 && K <= 8 && NST >= 2 && NST <= 8, "bad tile geometry");
  static constexpr int CH = 128 * K;            // message bytes per lane per stage
  static constexpr int TILE = 4096;             // 32 rows x 128 B
  static constexpr int STAGE_BYTES = K * TILE;
  static constexpr int WARP_BYTES = NST * STAGE_BYTES;  // tiles only; mbarriers live elsewhere
};

// ------------------------------------------------------------- ragged batches: cooperative row staging
// Messages of unlike lengths at unrelated addresses cannot be described by one tensor map, and one bulk copy per
// lane compiles to a 32-trip ELECT / R2UR / UBLKCP loop (38 % of the stall samples of that kernel).  Here the warp
// fills the same swizzled {128 B x 32 rows} tiles with plain cp.async (SASS LDGSTS.128): thread t moves the
// 16-byte piece (t & 7) of rows (t >> 3) + 4j, j = 0..7, so every instruction reads four whole 128-byte lines,
// and the eight row pointers / lengths a thread serves are fetched once, by shuffle, for the whole stream.
// The partial 16 bytes at a message's end are zero-filled by the copy's src-size; rows that have ended are skipped.
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, uint32_t src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

// Hashes one message per lane (same contract as digest_stream).  smem: Tiles<K,NST>::WARP_BYTES, 1024-byte aligned.
template <class A, int K, int NST>
__device__ __forceinline__ void digest_stream_rows(const uint8_t* __restrict__ msg, uint64_t len, uint32_t* h,
                                                   unsigned char* smem, int lane, uint32_t one) {
  using T = Tiles<K, NST>;
  constexpr int CH = T::CH;
  constexpr unsigned FULL = 0xffffffffu;
  const uint32_t s_base = smem_u32(smem);
  const uint32_t piece = (uint32_t)lane & 7u, r0 = (uint32_t)lane >> 3;
  const uint8_t* rp[8];
  uint64_t rl[8];
#pragma unroll
  for (int j = 0; j < 8; j++) {
    rp[j] = reinterpret_cast<const uint8_t*>(__shfl_sync(FULL, reinterpret_cast<unsigned long long>(msg), (int)r0 + 4 * j));
    rl[j] = __shfl_sync(FULL, (unsigned long long)len, (int)r0 + 4 * j);
  }
  const uint32_t nch = (uint32_t)(len / CH) + 1u;  // as in digest_stream: the last chunk is partial (possibly empty)
  const uint32_t last_bytes = (uint32_t)(len - (uint64_t)(nch - 1u) * CH);
  const uint32_t max_nch = __reduce_max_sync(FULL, nch);
  const uint32_t full_nch = __reduce_min_sync(FULL, nch) - 1u;  // chunks [0, full_nch) are whole in every row

  auto issue = [&](uint32_t c, int s) {
    const uint32_t d0 = s_base + (uint32_t)s * T::STAGE_BYTES;
    if (c < full_nch) {  // warp-uniform fast path: no bounds to check
      const uint64_t pos0 = (uint64_t)c * CH + piece * 16u;
#pragma unroll
      for (int j = 0; j < 8; j++) {
        const uint32_t r = r0 + 4u * (uint32_t)j;
        const uint32_t d = d0 + r * 128u + ((piece ^ (r & 7u)) << 4);
        const uint8_t* src = rp[j] + pos0;
#pragma unroll
        for (int k = 0; k < K; k++) cp_async16(d + (uint32_t)k * T::TILE, src + k * 128, 16u);
      }
    } else if (c < max_nch) {
      const uint64_t pos0 = (uint64_t)c * CH + piece * 16u;
#pragma unroll
      for (int j = 0; j < 8; j++) {
        const uint32_t r = r0 + 4u * (uint32_t)j;
        const uint32_t d = d0 + r * 128u + ((piece ^ (r & 7u)) << 4);
        if (pos0 < rl[j]) {
          const uint64_t rem64 = rl[j] - pos0;
          const int rem = rem64 > (uint64_t)(CH + 16) ? CH + 16 : (int)rem64;
          const uint8_t* src = rp[j] + pos0;
#pragma unroll
          for (int k = 0; k < K; k++) {
            const int rk = rem - 128 * k;
            if (rk > 0) cp_async16(d + (uint32_t)k * T::TILE, src + k * 128, rk >= 16 ? 16u : (uint32_t)rk);
          }
        }
      }
    }
    cp_async_commit();  // one group per call, empty or not, so that wait_group counts stages
  };

#pragma unroll
  for (int s = 0; s < NST; s++) issue((uint32_t)s, s);

  A::init(h);
  int s = 0;
  const uint32_t my = s_base + (uint32_t)lane * 128u;
  const uint32_t sw = (uint32_t)lane & 7u;
  auto load_block = [&](uint32_t src, uint32_t bb, uint32_t* w) {
    const uint32_t a = src + (bb >> 1) * T::TILE;
    const uint32_t q0 = (bb & 1u) * 4u;
#pragma unroll
    for (int q = 0; q < 4; q++) {
      uint4 v = lds128(a + (((q0 + (uint32_t)q) ^ sw) << 4));
      w[4 * q] = v.x; w[4 * q + 1] = v.y; w[4 * q + 2] = v.z; w[4 * q + 3] = v.w;
    }
  };
  for (uint32_t c = 0; c < max_nch; c++) {
    cp_async_wait<NST - 1>();
    __syncwarp();
    const uint32_t src = my + (uint32_t)s * T::STAGE_BYTES;
    // SHA-256 keeps ONE copy of its rounds: two would be 50 KB of code against a 32 KB instruction cache, and
    // warps in the tail loop beside warps in the body loop thrash it (measured 930 -> 800 GB/s).
    if (A::DLEN != 32 && c < full_nch) {  // every lane has CH / 64 whole blocks here: no padding logic in this loop
#pragma unroll 1
      for (uint32_t b = 0; b < (uint32_t)(CH / 64); b++) {
        uint32_t w[16];
        load_block(src, b, w);
        if (A::BE) {
#pragma unroll
          for (int i = 0; i < 16; i++) w[i] = bswap(w[i]);
        }
        A::compress(h, w, one);
      }
    } else if (c < nch) {
      const bool last = c == nch - 1u;
      const uint32_t bytes = last ? last_bytes : (uint32_t)CH;
      const uint32_t nb = bytes >> 6;
      const uint32_t r = bytes & 63u;
      const uint32_t nblk = nb + (last ? (r < 56 ? 1u : 2u) : 0u);
#pragma unroll 1
      for (uint32_t b = 0; b < nblk; b++) {
        uint32_t w[16];
        load_block(src, b < (uint32_t)(CH / 64) ? b : 0u, w);  // a second padding block reads nothing it keeps
        if (b >= nb) pad_block<A>(w, b - nb, r, len);
        if (A::BE) {
#pragma unroll
          for (int i = 0; i < 16; i++) w[i] = bswap(w[i]);
        }
        A::compress(h, w, one);
      }
    }
    __syncwarp();  // all lanes are done reading this stage before any lane's copies overwrite it
    issue(c + NST, s);
    if (++s == NST) s = 0;
  }
  cp_async_wait<0>();
}

// Per-warp pipeline state that survives across segments (stage cursor + one phase bit per stage).
struct PipeState {
  int stage = 0;
  uint32_t phase = 0;
};

// Runs chunks [c_begin, c_end) of the warp's 32 equal-length messages (rows row0 .. row0+31 of the
// tensor) through A::compress, continuing from chaining value h.  nch = len / CH + 1 is the chunk count
// of a whole message; its last chunk is partial (possibly empty) and carries the padding block(s).
// s_tiles: this warp's Tiles<K,NST>::WARP_BYTES (1024-byte aligned); s_bar: NST mbarriers (count 1).
template <class A, int K, int NST>
__device__ __forceinline__ void tile_segment(const void* tmap, uint32_t row0, uint64_t len, uint32_t c_begin,
                                             uint32_t c_end, uint32_t nch, uint32_t* h, uint32_t s_tiles,
                                             uint32_t s_bar, int lane, PipeState& ps, uint32_t one) {
  using G = Tiles<K, NST>;
  constexpr int CH = G::CH;
  auto issue = [&](uint32_t c, int s) {
    if (lane == 0) {
      const uint64_t pos = (uint64_t)c * CH;
      const uint64_t left = pos < len ? len - pos : 0;
      const uint32_t ntile = left >= (uint64_t)CH ? (uint32_t)K : (uint32_t)((left + 127) >> 7);
      const uint32_t bar = s_bar + 8 * s;
      mbar_arrive_expect_tx(bar, ntile * G::TILE);
#pragma unroll
      for (int j = 0; j < K; j++)
        if ((uint32_t)j < ntile) tma_tile_2d(s_tiles + (s * K + j) * G::TILE, tmap, (uint32_t)pos + 128u * j, row0, bar);
    }
  };
  {
    int s = ps.stage;
#pragma unroll
    for (int i = 0; i < NST; i++) {
      if (c_begin + i < c_end) issue(c_begin + i, s);
      if (++s == NST) s = 0;
    }
  }
  const uint32_t swz = (uint32_t)(lane & 7);
  for (uint32_t c = c_begin; c < c_end; c++) {
    const int s = ps.stage;
    mbar_wait(s_bar + 8 * s, (ps.phase >> s) & 1u);
    ps.phase ^= 1u << s;
    const bool last = c == nch - 1u;
    const uint32_t bytes = last ? (uint32_t)(len - (uint64_t)c * CH) : (uint32_t)CH;
    const uint32_t nb = bytes >> 6, r = bytes & 63u;
    const uint32_t nblk = nb + (last ? (r < 56 ? 1u : 2u) : 0u);
    const uint32_t lane_base = s_tiles + s * G::STAGE_BYTES + lane * 128;
    for (uint32_t b = 0; b < nblk; b++) {
      uint32_t w[16];
      const uint32_t bb = b < (uint32_t)(2 * K) ? b : 0u;  // padding block 1 past the stage: nothing to read
      const uint32_t a = lane_base + (bb >> 1) * G::TILE;
      const uint32_t q0 = (bb & 1u) * 4u;
#pragma unroll
      for (int q = 0; q < 4; q++) {
        uint4 v = lds128(a + (((q0 + q) ^ swz) << 4));
        w[4 * q] = v.x; w[4 * q + 1] = v.y; w[4 * q + 2] = v.z; w[4 * q + 3] = v.w;
      }
      if (b >= nb) pad_block<A>(w, b - nb, r, len);
      if (A::BE) {
#pragma unroll
        for (int i = 0; i < 16; i++) w[i] = bswap(w[i]);
      }
      A::compress(h, w, one);
    }
    __syncwarp();
    fence_proxy_async_smem();
    if (c + NST < c_end) issue(c + NST, s);
    if (++ps.stage == NST) ps.stage = 0;
  }
}

// The same stream for a warp that runs ALONE on its sub-partition (few large messages, MD5).  A lone MD5 chain is
// latency-bound: its step is a 3-deep dependent chain, 12.1 clocks, of which only 4 issue slots are used — so everything
// else a block needs (four LDS.128 and their swizzled addresses) is free IF it sits in the same basic block as a
// compression and does not feed it.  tile_segment loads a block, tests it for the tail, then compresses it: three
// basic blocks, all serial — 968 clocks per block measured against 807 for the bare compression
// (tools/microbench/chain.cu; ncu: 62 % of the samples are fixed-latency waits, 32 % issue).  Here the whole chunks of
// the message run through a loop with no branch in it but its own back edge: per trip two compressions, each with the
// loads of the block after next in front of it (two register sets, ping-pong, no moves).  The load that runs ahead
// crosses into the next stage at the end of a chunk, so that stage's barrier is waited for at the START of the chunk
// — which is why this wants NST >= 3: with two stages that chunk would only just have been requested.  The last,
// partial chunk with its padding block(s) goes through the general loop.
template <class A, int K, int NST>
__device__ __forceinline__ void tile_stream_lone(const void* tmap, uint32_t row0, uint64_t len, uint32_t nch, uint32_t* h,
                                                 uint32_t s_tiles, uint32_t s_bar, int lane, uint32_t one) {
  static_assert(NST >= 3, "the run-ahead load needs the next stage to have been requested a whole chunk earlier");
  using G = Tiles<K, NST>;
  constexpr int CH = G::CH;
  constexpr uint32_t RING = (uint32_t)(NST * K);  // tiles in the ring, consecutive in shared memory
  auto issue = [&](uint32_t c, int s) {
    if (lane == 0) {
      const uint64_t pos = (uint64_t)c * CH;
      const uint64_t left = pos < len ? len - pos : 0;
      const uint32_t ntile = left >= (uint64_t)CH ? (uint32_t)K : (uint32_t)((left + 127) >> 7);
      const uint32_t bar = s_bar + 8 * s;
      mbar_arrive_expect_tx(bar, ntile * G::TILE);
#pragma unroll
      for (int j = 0; j < K; j++)
        if ((uint32_t)j < ntile) tma_tile_2d(s_tiles + (s * K + j) * G::TILE, tmap, (uint32_t)pos + 128u * j, row0, bar);
    }
  };
#pragma unroll
  for (int i = 0; i < NST; i++)
    if ((uint32_t)i < nch) issue((uint32_t)i, i);
  const uint32_t swz = (uint32_t)(lane & 7);
  const uint32_t lane_base = s_tiles + (uint32_t)lane * 128u;
  // 16 words of half `half` (0 | 1) of ring tile `t`
  auto load_half = [&](uint32_t t, uint32_t half, uint32_t* w) {
    const uint32_t a = lane_base + t * G::TILE;
#pragma unroll
    for (int q = 0; q < 4; q++) {
      uint4 v = lds128(a + (((half * 4u + q) ^ swz) << 4));
      w[4 * q] = v.x; w[4 * q + 1] = v.y; w[4 * q + 2] = v.z; w[4 * q + 3] = v.w;
    }
    if (A::BE) {
#pragma unroll
      for (int i = 0; i < 16; i++) w[i] = bswap(w[i]);
    }
  };
  int stage = 0;
  uint32_t phase = 0;
  mbar_wait(s_bar, 0u);  // chunk 0 (whole, or the tail chunk of a short message)
  phase = 1u;
  uint32_t wa[16], wb[16];
  const uint32_t nfull = nch - 1u;  // whole chunks; chunk nch-1 is partial (possibly empty) and carries the padding
  if (nfull) load_half(0u, 0u, wa);
  for (uint32_t c = 0; c < nfull; c++) {
    int next = stage + 1;
    if (next == NST) next = 0;
    mbar_wait(s_bar + 8 * next, (phase >> next) & 1u);  // chunk c+1 <= nch-1 exists: the loop's last load reads its first block
    phase ^= 1u << next;
    uint32_t t = (uint32_t)stage * K;
#pragma unroll 1
    for (int i = 0; i < K; i++) {
      load_half(t, 1u, wb);
      A::compress(h, wa, one);
      t = t + 1u == RING ? 0u : t + 1u;
      load_half(t, 0u, wa);  // at i == K-1: the first block of the next stage (whose barrier has been waited for)
      A::compress(h, wb, one);
    }
    __syncwarp();
    fence_proxy_async_smem();
    if (c + NST < nch) issue(c + NST, stage);
    stage = next;
  }
  // tail chunk: its barrier has been waited for (above, or at the top when the message has no whole chunk)
  {
    const uint32_t bytes = (uint32_t)(len - (uint64_t)nfull * CH);
    const uint32_t nb = bytes >> 6, r = bytes & 63u;
    const uint32_t nblk = nb + (r < 56 ? 1u : 2u);
    const uint32_t t0 = (uint32_t)stage * K;
    for (uint32_t b = 0; b < nblk; b++) {
      uint32_t w[16];
      const uint32_t bb = b < (uint32_t)(2 * K) ? b : 0u;  // padding block 1 past the stage: nothing to read
      const uint32_t a = lane_base + (t0 + (bb >> 1)) * G::TILE;
#pragma unroll
      for (int q = 0; q < 4; q++) {
        uint4 v = lds128(a + ((((bb & 1u) * 4u + q) ^ swz) << 4));
        w[4 * q] = v.x; w[4 * q + 1] = v.y; w[4 * q + 2] = v.z; w[4 * q + 3] = v.w;
      }
      if (b >= nb) pad_block<A>(w, b - nb, r, len);
      if (A::BE) {
#pragma unroll
        for (int i = 0; i < 16; i++) w[i] = bswap(w[i]);
      }
      A::compress(h, w, one);
    }
  }
}

// ------------------------------------------------------- few, large messages: schedule off-loading
// When a GPU has fewer message groups than SM sub-partitions, every chain runs alone and the batch takes
// exactly one chain's latency, however many SMs idle.  A chain cannot be split, but its message schedule
// does not depend on the chaining value: SHA-256's W[16..63] (+K) is 34 % of the block's operations,
// SHA-1's W[16..79] 40 % of its ALU work.  So each group gets a second, "helper" warp on another
// sub-partition of the same SM that expands block b's schedule (byte swap, padding, K folded in) into a
// shared-memory ring while the chain warp runs the rounds of block b-1 on values it only has to LDS.
// Ring slots are handed over with named barriers (bar.arrive / bar.sync on 64 threads).
__device__ __forceinline__ void nb_sync(int id) { asm volatile("bar.sync %0, 64;" ::"r"(id) : "memory"); }
__device__ __forceinline__ void nb_arrive(int id) { asm volatile("bar.arrive %0, 64;" ::"r"(id) : "memory"); }
__device__ __forceinline__ void sts128(uint32_t addr, uint32_t x, uint32_t y, uint32_t z, uint32_t w) {
  asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(x), "r"(y), "r"(z), "r"(w) : "memory");
}

template <class A>
struct Assist;

template <>
struct Assist<Sha256> {
  static constexpr int WPB = 64;  // schedule words per block, K already added
  // w: the 16 big-endian message words.  Writes W[t]+K[t], t = 0..63, as 16 x 16 bytes per lane.
  // V = 0: every add an ALU-pipe IADD3 (11 ALU instructions per schedule word); V = 1: the adds are IMADs on the FMA
  // pipe (8 ALU + 4 FMA).  A warp alone on its sub-partition issues an instruction every 2.3 clocks while it feeds one
  // pipe and every 1.4 when it alternates between the two (tools/microbench/chain.cu), so the mix matters even
  // though nothing else competes for either pipe.
  // KH: the round constants of words t < KH are added here, the others by the chain warp (rounds<V, KH>).  Both warps
  // are issue-bound, so the adds go wherever they even out the pair (measured per block, tools/microbench/chain.cu:
  // 64 constants cost the helper 185 cycles and the chain warp 160).
  static constexpr int KH = 64;   // all in the helper: chain 1,630 cycles per block, helper 1,320 + ~300 of staging
  template <int V = 0, int KH_ = KH>
  __device__ static __forceinline__ void expand(uint32_t* w, uint32_t dst, uint32_t one) {
    constexpr uint32_t K[64] = {
        0x428a2f98, 0x71374491, 0xb5c0fbcf, 0xe9b5dba5, 0x3956c25b, 0x59f111f1, 0x923f82a4, 0xab1c5ed5,
        0xd807aa98, 0x12835b01, 0x243185be, 0x550c7dc3, 0x72be5d74, 0x80deb1fe, 0x9bdc06a7, 0xc19bf174,
        0xe49b69c1, 0xefbe4786, 0x0fc19dc6, 0x240ca1cc, 0x2de92c6f, 0x4a7484aa, 0x5cb0a9dc, 0x76f988da,
        0x983e5152, 0xa831c66d, 0xb00327c8, 0xbf597fc7, 0xc6e00bf3, 0xd5a79147, 0x06ca6351, 0x14292967,
        0x27b70a85, 0x2e1b2138, 0x4d2c6dfc, 0x53380d13, 0x650a7354, 0x766a0abb, 0x81c2c92e, 0x92722c85,
        0xa2bfe8a1, 0xa81a664b, 0xc24b8b70, 0xc76c51a3, 0xd192e819, 0xd6990624, 0xf40e3585, 0x106aa070,
        0x19a4c116, 0x1e376c08, 0x2748774c, 0x34b0bcb5, 0x391c0cb3, 0x4ed8aa4a, 0x5b9cca4f, 0x682e6ff3,
        0x748f82ee, 0x78a5636f, 0x84c87814, 0x8cc70208, 0x90befffa, 0xa4506ceb, 0xbef9a3f7, 0xc67178f2};
    uint32_t o[4];
#pragma unroll
    for (int t = 0; t < 64; t++) {
      if (t >= 16) {
        uint32_t x = w[(t - 15) & 15], y = w[(t - 2) & 15];
        uint32_t s0 = lop3<LUT_XOR3>(rotr(x, 7), rotr(x, 18), x >> 3);
        uint32_t s1 = lop3<LUT_XOR3>(rotr(y, 17), rotr(y, 19), y >> 10);
        if (V == 0) w[t & 15] = w[t & 15] + s0 + w[(t - 7) & 15] + s1;
        else w[t & 15] = add_fma(add_fma(w[t & 15], w[(t - 7) & 15], one), add_fma(s0, s1, one), one);
      }
      o[t & 3] = t >= KH_ ? w[t & 15] : V == 0 ? w[t & 15] + K[t] : add_fma(w[t & 15], K[t], one);
      if ((t & 3) == 3) sts128(dst + (t >> 2) * 512, o[0], o[1], o[2], o[3]);
    }
  }
  // The 64 rounds of one block on pre-expanded W+K read from `src` (this lane's 16-byte column).
  // The chain warp runs alone on its sub-partition, so its speed is its own issue rate: an instruction every ~2.3
  // clocks per pipe, ~1.4 when consecutive instructions alternate between the ALU and the FMA pipe (measured,
  // tools/microbench/chain.cu).  The rotates and booleans (10 per round) can only run on the ALU pipe; the variants
  // differ in where the round's adds go:
  //   V = 0  13 ALU + 2 FMA: three IADD3 (t1, e', a'), hh + wk and d + hh + wk as IMADs         (round 1: 29 clk/round)
  //   V = 1  11 ALU + 6 FMA: t1 and e' as IMAD pairs, a' = IADD3(t1, S0, maj)
  //   V = 2  10 ALU + 8 FMA: a' as an IMAD pair as well
  //   V = 3  10 ALU + 6 FMA: e' = d + t1 (one IMAD after t1, no separate d + hh + wk), a' as an IMAD pair
  //   V = 4  11 ALU + 4 FMA: as 3 but a' = IADD3
  // TRIPS: the 64 rounds as a loop of TRIPS trips (the working variables rotate with period 8, so a trip of 16 or 32
  // rounds ends where it began): 17 KB of code unrolled (TRIPS = 1), 8.7 KB with 2 trips, 4.5 KB with 4.  The chain
  // warp's loop, the helper's and — when MD5 of the same parts runs beside them, as a third warp or as its own kernel —
  // the MD5 loop share the SM's 32 KB instruction cache; with the rounds unrolled they add up to 37 KB and everything on
  // the SM slows down by a quarter (MD5 + SHA-256 over 4,096 x 16 MiB parts: 289 ms against 227 ms for SHA-256 alone,
  // although every warp had a sub-partition to itself).
  template <int V = 1, int KH_ = KH, int TRIPS = 1>
  __device__ static __forceinline__ void rounds(uint32_t* h, const uint4* src, uint32_t one) {
    constexpr uint32_t K[64] = {
        0x428a2f98, 0x71374491, 0xb5c0fbcf, 0xe9b5dba5, 0x3956c25b, 0x59f111f1, 0x923f82a4, 0xab1c5ed5,
        0xd807aa98, 0x12835b01, 0x243185be, 0x550c7dc3, 0x72be5d74, 0x80deb1fe, 0x9bdc06a7, 0xc19bf174,
        0xe49b69c1, 0xefbe4786, 0x0fc19dc6, 0x240ca1cc, 0x2de92c6f, 0x4a7484aa, 0x5cb0a9dc, 0x76f988da,
        0x983e5152, 0xa831c66d, 0xb00327c8, 0xbf597fc7, 0xc6e00bf3, 0xd5a79147, 0x06ca6351, 0x14292967,
        0x27b70a85, 0x2e1b2138, 0x4d2c6dfc, 0x53380d13, 0x650a7354, 0x766a0abb, 0x81c2c92e, 0x92722c85,
        0xa2bfe8a1, 0xa81a664b, 0xc24b8b70, 0xc76c51a3, 0xd192e819, 0xd6990624, 0xf40e3585, 0x106aa070,
        0x19a4c116, 0x1e376c08, 0x2748774c, 0x34b0bcb5, 0x391c0cb3, 0x4ed8aa4a, 0x5b9cca4f, 0x682e6ff3,
        0x748f82ee, 0x78a5636f, 0x84c87814, 0x8cc70208, 0x90befffa, 0xa4506ceb, 0xbef9a3f7, 0xc67178f2};
    static_assert(TRIPS == 1 || KH_ == 64, "the rolled loop has one body for all trips: the constants must come from the helper");
    static_assert(TRIPS == 1 || TRIPS == 2 || TRIPS == 4, "trips of 64, 32 or 16 rounds");
    uint32_t a = h[0], b = h[1], c = h[2], d = h[3], e = h[4], f = h[5], g = h[6], hh = h[7];
    // one round; t only matters (as a compile-time constant) when the chain warp adds the round constant itself
    auto round = [&](uint32_t wk, int t) {
      const uint32_t S1 = lop3<LUT_XOR3>(rotr(e, 6), rotr(e, 11), rotr(e, 25));
      const uint32_t ch = lop3<LUT_CH>(e, f, g);
      const uint32_t S0 = lop3<LUT_XOR3>(rotr(a, 2), rotr(a, 13), rotr(a, 22));
      const uint32_t mj = lop3<LUT_MAJ>(a, b, c);
      // hh + K[t] + w: nothing of this round's inputs; from word KH on the constant is added here (hh is four rounds old)
      const uint32_t p = t >= KH_ ? add_fma(add_fma(hh, K[t & 63], one), wk, one) : add_fma(hh, wk, one);
      uint32_t t1, en, an;
      if (V == 0) {
        const uint32_t dp = add_fma(d, p, one);
        t1 = p + ch + S1;
        en = dp + ch + S1;
        an = t1 + S0 + mj;
      } else if (V == 1 || V == 2) {
        const uint32_t dp = add_fma(d, p, one);
        t1 = add_fma(add_fma(p, ch, one), S1, one);
        en = add_fma(add_fma(dp, ch, one), S1, one);
        an = V == 1 ? t1 + S0 + mj : add_fma(t1, add_fma(S0, mj, one), one);
      } else {
        t1 = add_fma(add_fma(p, ch, one), S1, one);
        en = add_fma(d, t1, one);
        an = V == 4 ? t1 + S0 + mj : add_fma(t1, add_fma(S0, mj, one), one);
      }
      hh = g; g = f; f = e; e = en; d = c; c = b; b = a; a = an;
    };
    if (TRIPS > 1) {
      // The next trip's words are loaded into each half of the register window as soon as the current trip is done
      // with it, so no trip starts by waiting for its first LDS (that wait, ~35 clocks per trip, is most of what a
      // plain rolled loop costs: 1,833 clocks per block with 4 trips, 1,752 with this, 1,631 fully unrolled).
      constexpr int Q = 16 / TRIPS, HALF = Q / 2;  // 16-byte quads per trip
      uint4 v[Q];
#pragma unroll
      for (int q = 0; q < Q; q++) v[q] = src[q * 32];
#pragma unroll 1
      for (int trip = 0; trip < TRIPS; trip++) {
        const uint4* nx = src + ((trip + 1) % TRIPS) * (Q * 32);  // the last trip re-reads the first quads: harmless, unused
#pragma unroll
        for (int q = 0; q < HALF; q++) { round(v[q].x, 0); round(v[q].y, 0); round(v[q].z, 0); round(v[q].w, 0); }
#pragma unroll
        for (int q = 0; q < HALF; q++) v[q] = nx[q * 32];
#pragma unroll
        for (int q = HALF; q < Q; q++) { round(v[q].x, 0); round(v[q].y, 0); round(v[q].z, 0); round(v[q].w, 0); }
#pragma unroll
        for (int q = HALF; q < Q; q++) v[q] = nx[q * 32];
      }
    } else {
#pragma unroll
      for (int q = 0; q < 16; q++) {
        const uint4 v = src[q * 32];
        round(v.x, 4 * q); round(v.y, 4 * q + 1); round(v.z, 4 * q + 2); round(v.w, 4 * q + 3);
      }
    }
    h[0] += a; h[1] += b; h[2] += c; h[3] += d; h[4] += e; h[5] += f; h[6] += g; h[7] += hh;
  }
};

template <>
struct Assist<Sha1> {
  static constexpr int WPB = 80;
  // KH: the round constants of words t < KH are added here, the others by the chain warp (rounds<V, KH>).  Per block the
  // 80 constants cost the helper 219 cycles and the chain warp 62 (tools/microbench/chain.cu); with all of them in
  // the helper the pair ran at the helper's pace (ncu: the chain warp waited at the hand-over barrier), with all of
  // them in the chain warp at the chain's — the split evens the two out.
  static constexpr int KH = 80;   // all in the helper (a 32 / 48 split measured 2 % slower in the kernel than either end)
  template <int KH_ = KH>
  __device__ static __forceinline__ void expand(uint32_t* w, uint32_t dst, uint32_t one) {
    uint32_t o[4];
#pragma unroll
    for (int t = 0; t < 80; t++) {
      if (t >= 16)
        w[t & 15] = rotl(lop3<LUT_XOR3>(w[(t - 3) & 15], w[(t - 8) & 15], w[(t - 14) & 15]) ^ w[t & 15], 1);
      const uint32_t k = t < 20 ? 0x5a827999u : t < 40 ? 0x6ed9eba1u : t < 60 ? 0x8f1bbcdcu : 0xca62c1d6u;
      o[t & 3] = t >= KH_ ? w[t & 15] : add_fma(w[t & 15], k, one);
      if ((t & 3) == 3) sts128(dst + (t >> 2) * 512, o[0], o[1], o[2], o[3]);
    }
  }
  // V = 0: x = IADD3(e, wk, f) — 4 ALU instructions per round; V = 1: x = (e + wk) + f as two IMADs — 3 ALU + 2 FMA
  // (e + wk is off the chain; the second IMAD is on it).  See Assist<Sha256>::rounds for why the mix matters.
  // Measured per block: 702 / 780.  (rotl(b, 30) as IMAD.HI + IMAD — hi and lo word of b * 2^30 — to take the fourth
  // instruction off the ALU pipe: 965; IMAD.HI is no full-rate instruction.)
  template <int V = 0, int KH_ = KH>
  __device__ static __forceinline__ void rounds(uint32_t* h, const uint4* src, uint32_t one) {
    uint32_t a = h[0], b = h[1], c = h[2], d = h[3], e = h[4];
#pragma unroll
    for (int q = 0; q < 20; q++) {
      const uint4 v = src[q * 32];
      const uint32_t wk[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int j = 0; j < 4; j++) {
        const int t = 4 * q + j;
        uint32_t f = t < 20 ? lop3<LUT_CH>(b, c, d) : t < 40 ? lop3<LUT_XOR3>(b, c, d)
                   : t < 60 ? lop3<LUT_MAJ>(b, c, d) : lop3<LUT_XOR3>(b, c, d);
        const uint32_t k = t < 20 ? 0x5a827999u : t < 40 ? 0x6ed9eba1u : t < 60 ? 0x8f1bbcdcu : 0xca62c1d6u;
        const uint32_t ek = t >= KH_ ? add_fma(e, k, one) : e;  // e is four rounds old: an IMAD on the otherwise idle FMA pipe
        uint32_t x = V == 0 ? ek + wk[j] + f : add_fma(add_fma(ek, wk[j], one), f, one);
        uint32_t tmp = rotl(a, 5) + x;
        e = d; d = c; c = rotl(b, 30); b = a; a = tmp;
      }
    }
    h[0] += a; h[1] += b; h[2] += c; h[3] += d; h[4] += e;
  }
};

// Helper warp: streams the group's message through the TMA tiles (as tile_segment does) and, instead of
// compressing, expands every block (padding blocks included) into schedule-ring slot (block % SS).
// bar0: first of the 2*SS named barriers this chain / helper pair hands slots over with (a CTA may hold two pairs).
// BPS: blocks per ring slot — a slot is handed over when it holds BPS expanded blocks (or the message's last one), so
// the two bar.sync / bar.arrive a hand-over costs each warp are paid once per BPS blocks.
// SHARED: the tiles belong to another warp (tile_consume_shared<..., ISSUER>), which issues the TMA loads; this warp only
// waits for a stage's `full` mbarrier and, when it is done with the stage, arrives on its `empty` mbarrier (s_empty).
// The helper is the warp with the least slack of the MD5 + SHA-256 trio, so it is spared the issuing.
template <class A, int K, int NST, int SS, int BPS = 1, bool SHARED = false>
__device__ __forceinline__ void assist_produce(const void* tmap, uint32_t row0, uint64_t len, uint32_t nch,
                                               uint32_t s_tiles, uint32_t s_bar, uint32_t s_ring, int lane,
                                               uint32_t one, int bar0 = 1, uint32_t s_empty = 0) {
  using G = Tiles<K, NST>;
  constexpr int CH = G::CH;
  constexpr uint32_t SLOT = Assist<A>::WPB * 128;
  auto issue = [&](uint32_t c, int s) {
    if (lane == 0) {
      const uint64_t pos = (uint64_t)c * CH;
      const uint64_t left = pos < len ? len - pos : 0;
      const uint32_t ntile = left >= (uint64_t)CH ? (uint32_t)K : (uint32_t)((left + 127) >> 7);
      const uint32_t bar = s_bar + 8 * s;
      mbar_arrive_expect_tx(bar, ntile * G::TILE);
#pragma unroll
      for (int j = 0; j < K; j++)
        if ((uint32_t)j < ntile) tma_tile_2d(s_tiles + (s * K + j) * G::TILE, tmap, (uint32_t)pos + 128u * j, row0, bar);
    }
  };
  if (!SHARED) {
#pragma unroll
    for (int i = 0; i < NST; i++)
      if ((uint32_t)i < nch) issue((uint32_t)i, i);
  }
  const uint32_t swz = (uint32_t)(lane & 7);
  int stage = 0, slot = 0, inslot = 0;
  uint32_t phase = 0, produced = 0;  // produced: slots handed over so far
  for (uint32_t c = 0; c < nch; c++) {
    mbar_wait(s_bar + 8 * stage, (phase >> stage) & 1u);
    phase ^= 1u << stage;
    const bool last = c == nch - 1u;
    const uint32_t bytes = last ? (uint32_t)(len - (uint64_t)c * CH) : (uint32_t)CH;
    const uint32_t nb = bytes >> 6, r = bytes & 63u;
    const uint32_t nblk = nb + (last ? (r < 56 ? 1u : 2u) : 0u);
    const uint32_t lane_base = s_tiles + stage * G::STAGE_BYTES + lane * 128;
    for (uint32_t b = 0; b < nblk; b++) {
      uint32_t w[16];
      const uint32_t bb = b < (uint32_t)(2 * K) ? b : 0u;
      const uint32_t a = lane_base + (bb >> 1) * G::TILE;
      const uint32_t q0 = (bb & 1u) * 4u;
#pragma unroll
      for (int q = 0; q < 4; q++) {
        uint4 v = lds128(a + (((q0 + q) ^ swz) << 4));
        w[4 * q] = v.x; w[4 * q + 1] = v.y; w[4 * q + 2] = v.z; w[4 * q + 3] = v.w;
      }
      if (b >= nb) pad_block<A>(w, b - nb, r, len);
#pragma unroll
      for (int i = 0; i < 16; i++) w[i] = bswap(w[i]);
      if (inslot == 0 && produced >= (uint32_t)SS) nb_sync(bar0 + SS + slot);  // wait until the chain warp has drained this slot
      Assist<A>::expand(w, s_ring + (slot * BPS + inslot) * SLOT + lane * 16, one);
      if (++inslot == BPS || (last && b == nblk - 1u)) {  // slot full, or the message's last block: hand it over
        nb_arrive(bar0 + slot);
        produced++;
        inslot = 0;
        if (++slot == SS) slot = 0;
      }
    }
    __syncwarp();
    fence_proxy_async_smem();
    if (c + NST < nch) {
      if (SHARED) { if (lane == 0) mbar_arrive(s_empty + 8 * stage); }
      else issue(c + NST, stage);
    }
    if (++stage == NST) stage = 0;
  }
}

// The warp that owns the tiles of the one-launch MD5 + SHA-256 kernel: it issues the TMA loads, compresses every block
// of the group's messages with digest A (MD5) as the stages fill, and refills a stage once it AND the second reader of
// the tiles (assist_produce<..., SHARED>, the SHA-256 helper) have arrived on the stage's `empty` mbarrier (count 2) —
// so the parts are read from HBM once.  The MD5 chain needs half the time per block the SHA-256 pair does, so this
// warp has the slack for the issuing and the waiting.
template <class A, int K, int NST>
__device__ __forceinline__ void tile_consume_shared(const void* tmap, uint32_t row0, uint64_t len, uint32_t nch, uint32_t* h,
                                                    uint32_t s_tiles, uint32_t s_full, uint32_t s_empty, int lane, uint32_t one) {
  using G = Tiles<K, NST>;
  constexpr int CH = G::CH;
  auto issue = [&](uint32_t c, int s) {
    if (lane == 0) {
      const uint64_t pos = (uint64_t)c * CH;
      const uint64_t left = pos < len ? len - pos : 0;
      const uint32_t ntile = left >= (uint64_t)CH ? (uint32_t)K : (uint32_t)((left + 127) >> 7);
      const uint32_t bar = s_full + 8 * s;
      mbar_arrive_expect_tx(bar, ntile * G::TILE);
#pragma unroll
      for (int j = 0; j < K; j++)
        if ((uint32_t)j < ntile) tma_tile_2d(s_tiles + (s * K + j) * G::TILE, tmap, (uint32_t)pos + 128u * j, row0, bar);
    }
  };
#pragma unroll
  for (int i = 0; i < NST; i++)
    if ((uint32_t)i < nch) issue((uint32_t)i, i);
  const uint32_t swz = (uint32_t)(lane & 7);
  int stage = 0;
  uint32_t phase = 0, ephase = 0;
  for (uint32_t c = 0; c < nch; c++) {
    mbar_wait(s_full + 8 * stage, (phase >> stage) & 1u);
    phase ^= 1u << stage;
    const bool last = c == nch - 1u;
    const uint32_t bytes = last ? (uint32_t)(len - (uint64_t)c * CH) : (uint32_t)CH;
    const uint32_t nb = bytes >> 6, r = bytes & 63u;
    const uint32_t nblk = nb + (last ? (r < 56 ? 1u : 2u) : 0u);
    const uint32_t lane_base = s_tiles + stage * G::STAGE_BYTES + lane * 128;
#pragma unroll 1
    for (uint32_t b = 0; b < nblk; b++) {
      uint32_t w[16];
      const uint32_t bb = b < (uint32_t)(2 * K) ? b : 0u;  // padding block 1 past the stage: nothing to read
      const uint32_t a = lane_base + (bb >> 1) * G::TILE;
      const uint32_t q0 = (bb & 1u) * 4u;
#pragma unroll
      for (int q = 0; q < 4; q++) {
        uint4 v = lds128(a + (((q0 + q) ^ swz) << 4));
        w[4 * q] = v.x; w[4 * q + 1] = v.y; w[4 * q + 2] = v.z; w[4 * q + 3] = v.w;
      }
      if (b >= nb) pad_block<A>(w, b - nb, r, len);
      if (A::BE) {
#pragma unroll
        for (int i = 0; i < 16; i++) w[i] = bswap(w[i]);
      }
      A::compress(h, w, one);
    }
    __syncwarp();
    fence_proxy_async_smem();
    if (c + NST < nch) {
      if (lane == 0) {
        mbar_arrive(s_empty + 8 * stage);
        mbar_wait(s_empty + 8 * stage, (ephase >> stage) & 1u);  // the other reader is done with this stage too
      }
      ephase ^= 1u << stage;
      issue(c + NST, stage);
    }
    if (++stage == NST) stage = 0;
  }
}

// Chain warp: consumes `total` expanded blocks.
template <class A, int SS, int TRIPS = 1, int BPS = 1>
__device__ __forceinline__ void assist_consume(uint32_t total, uint32_t* h, const unsigned char* ring, int lane,
                                               uint32_t one, int bar0 = 1) {
  constexpr uint32_t SLOT = Assist<A>::WPB * 128;
  const uint32_t nslots = (total + BPS - 1) / BPS;
  int slot = 0;
  uint32_t done = 0;
  for (uint32_t i = 0; i < nslots; i++) {
    nb_sync(bar0 + slot);
    const uint32_t nb = total - done < (uint32_t)BPS ? total - done : (uint32_t)BPS;
#pragma unroll 1  // one copy of the rounds: two would not fit the 32 KB instruction cache beside the helper's loop
    for (uint32_t j = 0; j < nb; j++) {
      const uint4* src = reinterpret_cast<const uint4*>(ring + (slot * BPS + j) * SLOT) + lane;
      if constexpr (TRIPS == 1) Assist<A>::rounds(h, src, one);
      else Assist<A>::template rounds<1, Assist<A>::KH, TRIPS>(h, src, one);
    }
    done += nb;
    if (i + SS < nslots) nb_arrive(bar0 + SS + slot);  // hand the slot back (only if the helper will reuse it)
    if (++slot == SS) slot = 0;
  }
}

// Writes the digest in its standard byte order (SHA: big-endian words; MD5: little-endian words).
template <class A>
__device__ __forceinline__ void store_digest(uint8_t* out, const uint32_t* h) {
  uint32_t* o = reinterpret_cast<uint32_t*>(out);  // 4-byte aligned: out + i*{20,16,32}
#pragma unroll
  for (int i = 0; i < A::NS; i++) o[i] = A::BE ? bswap(h[i]) : h[i];
}

}  // namespace th
