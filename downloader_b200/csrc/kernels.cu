// kernels.cu — sm_100a kernels and the launch-level C ABI (include/tritonhash.h, last section).
//
//   digest_rows_kernel<A,K,NST> / digest_kernel<A,CH,NST>  one warp = one CTA = 32 messages of any lengths;
//                            th::digest_stream_rows (default) / th::digest_stream do the work.
//                            SHA-1 adds the fused compare -> verified bitfield ([dep] webtorrent
//                            `hash === _hashes[index]` -> bitfield.set(index); lib/download.js:64-66).
//   etag_kernel              one warp per multipart object, md5 over its part digests, message words
//                            exchanged by warp shuffle (S3 ETag returned to [dep] minio
//                            completeMultipartUpload; lib/upload.js:45).
//   synth_kernel             SURVEY §8d counter-based generator (benchmarks / tests).
#include <cstdint>
#include <atomic>
#include <cstdio>
#include <string>
#include <cuda.h>  // CUtensorMap types only; the encoder is fetched with cudaGetDriverEntryPoint (no -lcuda)
#include <cuda_runtime.h>

#include "../../include/tritonhash.h"
#include "digest_device.cuh"

namespace th {

struct DigestArgs {
  const uint8_t* base;
  const uint64_t* offsets;  // nullable -> i * stride
  const uint64_t* lengths;  // nullable -> len
  const uint32_t* order;    // nullable -> identity
  uint64_t stride, len;
  uint32_t n;
  uint8_t* out;
  const uint8_t* expected;  // SHA-1 only, nullable
  uint32_t* ok_bits;        // SHA-1 only, nullable
  uint32_t one;             // the value 1, opaque to the compiler (see add_fma)
};

// Which message a lane hashes: slot -> (order) -> offset / length.
__device__ __forceinline__ bool lane_message(const DigestArgs& p, int lane, uint32_t* idx, uint64_t* off, uint64_t* len) {
  const uint32_t slot = blockIdx.x * 32u + (uint32_t)lane;
  const bool valid = slot < p.n;
  *idx = slot;
  if (valid && p.order) *idx = p.order[slot];
  *off = 0;
  *len = 0;
  if (valid) {
    *off = p.offsets ? p.offsets[*idx] : (uint64_t)*idx * p.stride;
    *len = p.lengths ? p.lengths[*idx] : p.len;
  }
  return valid;
}

// Digest store + (SHA-1) fused piece verification for the per-lane kernels.
template <class A>
__device__ __forceinline__ void lane_finish(const DigestArgs& p, int lane, bool valid, uint32_t idx, const uint32_t* h) {
  if (valid) store_digest<A>(p.out + (size_t)idx * A::DLEN, h);
  if (A::DLEN == 20 && p.expected && p.ok_bits) {
    bool ok = false;
    if (valid) {
      const uint32_t* e = reinterpret_cast<const uint32_t*>(p.expected + (size_t)idx * 20);
      ok = true;
#pragma unroll
      for (int i = 0; i < 5; i++) ok &= (__ldg(e + i) == bswap(h[i]));
    }
    if (p.order) {
      if (ok) atomicOr(p.ok_bits + (idx >> 5), 1u << (idx & 31));
    } else {
      const uint32_t word = __ballot_sync(0xffffffffu, ok);
      if (lane == 0) p.ok_bits[blockIdx.x] = word;
    }
  }
}

template <class A, int CH, int NST>
__global__ void __launch_bounds__(32) digest_kernel(const DigestArgs p) {
  extern __shared__ __align__(1024) unsigned char smem[];
  const int lane = threadIdx.x;
  uint32_t idx;
  uint64_t off, len;
  const bool valid = lane_message(p, lane, &idx, &off, &len);
  uint32_t h[A::NS];
  digest_stream<A, CH, NST>(p.base + off, len, h, smem, lane, p.one);
  lane_finish<A>(p, lane, valid, idx, h);
}

// Same contract, staged by cooperative cp.async into swizzled row tiles (digest_stream_rows): the ragged default.
template <class A, int K, int NST>
__global__ void __launch_bounds__(32) digest_rows_kernel(const DigestArgs p) {
  extern __shared__ __align__(1024) unsigned char smem[];
  const int lane = threadIdx.x;
  uint32_t idx;
  uint64_t off, len;
  const bool valid = lane_message(p, lane, &idx, &off, &len);
  uint32_t h[A::NS];
  digest_stream_rows<A, K, NST>(p.base + off, len, h, smem, lane, p.one);
  lane_finish<A>(p, lane, valid, idx, h);
}

// --------------------------------------------------------------- equal-length batches: tensor-map path
struct TileArgs {
  uint64_t len;
  uint32_t n;
  uint32_t segc;            // sliced kernel: chunks per segment
  uint32_t nseg;            // sliced kernel: segments per message
  uint8_t* out;
  uint8_t* out_md5;         // digest_parts_kernel: the MD5 digests (out holds the SHA-256 ones)
  const uint8_t* expected;  // SHA-1 only, nullable
  uint32_t* ok_bits;        // SHA-1 only, nullable
  uint32_t one;             // the value 1, opaque to the compiler (see add_fma)
  uint32_t salt;            // small grids: CTA b works on group (b + gridDim.x - salt) % gridDim.x, see spread_group
};

// A kernel with only a few CTAs always gets the same first SMs, so several such kernels running at once
// (tickets of different ring slots, MD5 beside SHA-256) pile their long chains onto the same sub-partitions
// while most of the chip idles: measured 1.05 s -> 1.45..1.9 s for eight concurrent 64 MiB-part submits.
// Small grids are therefore padded to one CTA per SM and rotated by a per-launch salt; CTAs without a group
// exit at once.  Returns false for such a CTA.
__device__ __forceinline__ bool spread_group(const TileArgs& p, uint32_t* group) {
  const uint32_t groups = (p.n + 31u) / 32u;
  uint32_t g = blockIdx.x;
  if (gridDim.x > groups) {
    g = blockIdx.x + gridDim.x - p.salt;
    if (g >= gridDim.x) g -= gridDim.x;
  }
  *group = g;
  return g < groups;
}

template <class A>
__device__ __forceinline__ void finish_group(const TileArgs& p, uint32_t group, int lane, const uint32_t* h) {
  const uint32_t idx = group * 32u + lane;
  const bool valid = idx < p.n;
  if (valid) store_digest<A>(p.out + (size_t)idx * A::DLEN, h);
  if (A::DLEN == 20 && p.expected && p.ok_bits) {  // fused piece verification -> bitfield word
    bool ok = false;
    if (valid) {
      const uint32_t* e = reinterpret_cast<const uint32_t*>(p.expected + (size_t)idx * 20);
      ok = true;
#pragma unroll
      for (int i = 0; i < 5; i++) ok &= (__ldg(e + i) == bswap(h[i]));
    }
    const uint32_t word = __ballot_sync(0xffffffffu, ok);
    if (lane == 0) p.ok_bits[group] = word;
  }
}

// One warp = one CTA = 32 consecutive messages; used when messages are too few to fill the chip
// (every group gets its own warp slot, and MD5 / SHA-256 kernels of one submit co-reside on an SM).
// LONE: the warp has its sub-partition to itself (few large messages); the message then runs through
// tile_stream_lone, whose loop overlaps a block's loads with the previous block's compression.
template <class A, int K, int NST, bool LONE = false>
__global__ void __launch_bounds__(32) digest_tile_kernel(const __grid_constant__ CUtensorMap tmap, const TileArgs p) {
  extern __shared__ __align__(1024) unsigned char smem[];
  using G = Tiles<K, NST>;
  const int lane = threadIdx.x;
  uint32_t group;
  if (!spread_group(p, &group)) return;
  const uint32_t s_tiles = smem_u32(smem);
  const uint32_t s_bar = s_tiles + G::WARP_BYTES;
  if (lane == 0) {
    prefetch_tmap(&tmap);
#pragma unroll
    for (int s = 0; s < NST; s++) mbar_init(s_bar + 8 * s, 1);
    mbar_init_fence();
  }
  __syncwarp();
  const uint32_t nch = (uint32_t)(p.len / G::CH) + 1u;
  uint32_t h[A::NS];
  A::init(h);
  if constexpr (LONE) {
    tile_stream_lone<A, K, NST>(&tmap, group * 32u, p.len, nch, h, s_tiles, s_bar, lane, p.one);
  } else {
    PipeState ps;
    tile_segment<A, K, NST>(&tmap, group * 32u, p.len, 0u, nch, nch, h, s_tiles, s_bar, lane, ps, p.one);
  }
  finish_group<A>(p, group, lane, h);
}

// Few, large messages (fewer groups than SM sub-partitions): chain warp + schedule helper warp per group.
// SS ring slots of BPS expanded blocks each.
template <class A, int K, int NST, int SS, int BPS = 1>
__global__ void __launch_bounds__(64) digest_assist_kernel(const __grid_constant__ CUtensorMap tmap, const TileArgs p) {
  extern __shared__ __align__(1024) unsigned char smem[];
  using G = Tiles<K, NST>;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t group;
  if (!spread_group(p, &group)) return;
  const uint32_t s_tiles = smem_u32(smem);
  const uint32_t s_ring = s_tiles + G::WARP_BYTES;
  const uint32_t s_bar = s_ring + SS * BPS * Assist<A>::WPB * 128;
  if (threadIdx.x == 32) {
    prefetch_tmap(&tmap);
#pragma unroll
    for (int s = 0; s < NST; s++) mbar_init(s_bar + 8 * s, 1);
    mbar_init_fence();
  }
  __syncthreads();
  const uint32_t nch = (uint32_t)(p.len / G::CH) + 1u;
  const uint32_t r = (uint32_t)(p.len & 63u);
  const uint32_t total = (uint32_t)(p.len >> 6) + (r < 56 ? 1u : 2u);
  if (warp == 1) {
    assist_produce<A, K, NST, SS, BPS>(&tmap, group * 32u, p.len, nch, s_tiles, s_bar, s_ring, lane, p.one);
  } else {
    uint32_t h[A::NS];
    A::init(h);
    assist_consume<A, SS, 1, BPS>(total, h, smem + G::WARP_BYTES, lane, p.one);
    finish_group<A>(p, group, lane, h);
  }
}

// MD5 AND SHA-256 of few large messages — the digests an S3 part needs (lib/upload.js:45 -> [dep] minio getHashSummer)
// — in ONE launch that reads the parts once and fixes which sub-partition every warp runs on.  Launched as two kernels
// (MD5 tile kernel + SHA-256 chain/helper kernel) config 3 took 246..294 ms from run to run against 227 ms for SHA-256
// alone, and read every part twice.  Warps of one CTA are dealt to the four sub-partitions in order, so the roles are
// fixed by warp index:
//   GP = 1 (up to one group per SM):  warp 0 SHA-256 chain, 1 schedule helper, 2 MD5 chain            (3 sub-partitions)
//   GP = 2 (up to two groups per SM): warps 0,1 the two SHA-256 chains, 2,3 their helpers, 4,5 exit at once,
//                                     6,7 the two MD5 chains — which share sub-partitions 2,3 with the helpers, the
//                                     warps with issue slots to spare, never with a chain warp.
// The MD5 warp owns the tiles and issues the TMA loads (tile_consume_shared); the helper reads the same tiles: both wait
// on a stage's `full` mbarrier, and the stage is refilled only after both have arrived on its `empty` mbarrier.  The SHA-256 rounds run as TRIPS trips so
// that chain loop + helper loop + MD5 loop fit the 32 KB instruction cache (see Assist<Sha256>::rounds).
template <int K, int NST, int SS, int BPS, int GP, int TRIPS>
__global__ void __launch_bounds__(GP == 1 ? 96 : 256) digest_parts_kernel(const __grid_constant__ CUtensorMap tmap, const TileArgs p) {
  extern __shared__ __align__(1024) unsigned char smem[];
  using G = Tiles<K, NST>;
  constexpr uint32_t RING_BYTES = SS * BPS * Assist<Sha256>::WPB * 128;
  constexpr uint32_t PER_GROUP = G::WARP_BYTES + RING_BYTES;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  // role and which of the CTA's groups this warp serves
  int role, g;  // role 0 chain, 1 helper, 2 md5, 3 none
  if (GP == 1) { role = warp; g = 0; }
  else { role = warp < 2 ? 0 : warp < 4 ? 1 : warp < 6 ? 3 : 2; g = warp & 1; }
  const uint32_t groups = (p.n + 31u) / 32u;
  uint32_t slot = blockIdx.x;  // CTA -> slot of GP groups, rotated by the per-launch salt like spread_group
  const uint32_t slots = (groups + GP - 1) / GP;
  if (gridDim.x > slots) {
    slot = blockIdx.x + gridDim.x - p.salt;
    if (slot >= gridDim.x) slot -= gridDim.x;
  }
  const uint32_t group = slot * GP + (uint32_t)g;
  const bool active = slot < slots && group < groups && role != 3;
  unsigned char* base = smem + (size_t)g * PER_GROUP;
  const uint32_t s_tiles = smem_u32(base);
  const uint32_t s_ring = s_tiles + G::WARP_BYTES;
  const uint32_t s_full = smem_u32(smem) + GP * PER_GROUP + (uint32_t)g * 64u;  // per group: NST full + NST empty mbarriers
  const uint32_t s_empty = s_full + 8 * NST;
  if (active && lane == 0 && role == 2) {
    prefetch_tmap(&tmap);
#pragma unroll
    for (int s = 0; s < NST; s++) {
      mbar_init(s_full + 8 * s, 1);
      mbar_init(s_empty + 8 * s, 2);  // helper + MD5 warp
    }
    mbar_init_fence();
  }
  __syncthreads();
  if (!active) return;
  const uint32_t r = (uint32_t)(p.len & 63u);
  const uint32_t total = (uint32_t)(p.len >> 6) + (r < 56 ? 1u : 2u);
  const uint32_t nch = (uint32_t)(p.len / G::CH) + 1u;
  const int bar0 = 1 + g * 2 * SS;
  if (role == 1) {
    assist_produce<Sha256, K, NST, SS, BPS, true>(&tmap, group * 32u, p.len, nch, s_tiles, s_full, s_ring, lane, p.one, bar0, s_empty);
  } else if (role == 0) {
    uint32_t h[Sha256::NS];
    Sha256::init(h);
    assist_consume<Sha256, SS, TRIPS, BPS>(total, h, base + G::WARP_BYTES, lane, p.one, bar0);  // rolled rounds: see Assist<Sha256>::rounds
    finish_group<Sha256>(p, group, lane, h);
  } else {
    uint32_t h[Md5::NS];
    Md5::init(h);
    tile_consume_shared<Md5, K, NST>(&tmap, group * 32u, p.len, nch, h, s_tiles, s_full, s_empty, lane, p.one);
    const uint32_t idx = group * 32u + lane;
    if (idx < p.n) store_digest<Md5>(p.out_md5 + (size_t)idx * Md5::DLEN, h);
  }
}

// Persistent, time-sliced: one CTA per SM, W worker warps (a multiple of 4, so every SM sub-partition
// holds the same number), G = this SM's share of the message groups.  A chain cannot be split across
// threads, but it can be handed from warp to warp between blocks: its whole state is the chaining value.
// So each group's message is cut into `nseg` segments and unit k = (segment k / G, group k % G) is run
// by worker k % W.  With G > W the G chains advance round-robin at the same rate and all four
// sub-partitions stay equally loaded until the end, instead of 4,4,3,3 warps finishing at different
// times.  Chaining values travel through shared memory; done[g] counts finished segments.
template <class A, int K, int NST>
__global__ void __launch_bounds__(A::SLICED_BOUND, 1) digest_sliced_kernel(const __grid_constant__ CUtensorMap tmap,
                                                               const TileArgs p) {
  extern __shared__ __align__(1024) unsigned char smem[];
  using G_ = Tiles<K, NST>;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t W = blockDim.x >> 5;
  const uint32_t groups_total = (p.n + 31u) / 32u;
  const uint32_t g0 = (uint32_t)(((uint64_t)blockIdx.x * groups_total) / gridDim.x);
  const uint32_t g1 = (uint32_t)(((uint64_t)(blockIdx.x + 1) * groups_total) / gridDim.x);
  const uint32_t G = g1 - g0;
  const uint32_t s_tiles = smem_u32(smem) + warp * G_::WARP_BYTES;
  const uint32_t s_bar = smem_u32(smem) + W * G_::WARP_BYTES + warp * (8 * NST);
  // after the tiles and barriers: done[G] (u32, padded to 128 B), then state[G][NS][32]
  unsigned char* tail = smem + W * G_::WARP_BYTES + ((W * 8 * NST + 127) & ~127u);
  volatile uint32_t* done = reinterpret_cast<volatile uint32_t*>(tail);
  uint32_t* state = reinterpret_cast<uint32_t*>(tail + ((G * 4 + 127) & ~127u));

  if (lane == 0) {
    if (warp == 0) prefetch_tmap(&tmap);
#pragma unroll
    for (int s = 0; s < NST; s++) mbar_init(s_bar + 8 * s, 1);
    mbar_init_fence();
  }
  const uint32_t S = p.nseg;
  if (S > 1)
    for (uint32_t g = threadIdx.x; g < G; g += blockDim.x) done[g] = 0;
  __syncthreads();

  const uint32_t nch = (uint32_t)(p.len / G_::CH) + 1u;
  PipeState ps;
  const uint32_t units = G * S;
  for (uint32_t k = warp; k < units; k += W) {
    const uint32_t g = k % G, seg = k / G;
    uint32_t h[A::NS];
    if (seg == 0) {
      A::init(h);
    } else {
      if (lane == 0) {
        while (done[g] < seg) __nanosleep(64);
      }
      __syncwarp();
      __threadfence_block();
#pragma unroll
      for (int i = 0; i < A::NS; i++) h[i] = state[(g * A::NS + i) * 32 + lane];
    }
    const uint32_t c_begin = seg * p.segc;
    const uint32_t c_end = min(nch, c_begin + p.segc);
    tile_segment<A, K, NST>(&tmap, (g0 + g) * 32u, p.len, c_begin, c_end, nch, h, s_tiles, s_bar, lane, ps, p.one);
    if (seg == S - 1) {
      finish_group<A>(p, g0 + g, lane, h);
    } else {
#pragma unroll
      for (int i = 0; i < A::NS; i++) state[(g * A::NS + i) * 32 + lane] = h[i];
      __threadfence_block();
      __syncwarp();
      if (lane == 0) done[g] = seg + 1;
    }
  }
}

// ------------------------------------------------------------------------------------ ETag combine
// One warp per object.  The message is the object's part digests back to back (N*16 bytes, so the
// tail is 0/16/32/48 bytes and there is always exactly one padding block).  Each iteration the warp
// loads a 128-byte window (two MD5 blocks) with one coalesced 4-byte load per lane; every lane runs
// the same chain and fetches message word g from lane g (block A) / 16+g (block B) by shuffle.
struct Md5Shfl {
  template <int I>
  __device__ static __forceinline__ void step(uint32_t& a, uint32_t b, uint32_t c, uint32_t d, uint32_t mine,
                                              int half) {
    constexpr int R = I >> 4;
    constexpr int G = R == 0 ? I : R == 1 ? (5 * I + 1) & 15 : R == 2 ? (3 * I + 5) & 15 : (7 * I) & 15;
    uint32_t m[16];
    m[G] = __shfl_sync(0xffffffffu, mine, half * 16 + G);
    Md5::step<I>(a, b, c, d, m, 1u);
  }
  template <int I>
  __device__ static __forceinline__ void four(uint32_t& a, uint32_t& b, uint32_t& c, uint32_t& d, uint32_t mine,
                                              int half) {
    step<I>(a, b, c, d, mine, half);
    step<I + 1>(d, a, b, c, mine, half);
    step<I + 2>(c, d, a, b, mine, half);
    step<I + 3>(b, c, d, a, mine, half);
  }
  __device__ static __forceinline__ void block(uint32_t* h, uint32_t mine, int half) {
    uint32_t a = h[0], b = h[1], c = h[2], d = h[3];
    four<0>(a, b, c, d, mine, half);  four<4>(a, b, c, d, mine, half);  four<8>(a, b, c, d, mine, half);
    four<12>(a, b, c, d, mine, half); four<16>(a, b, c, d, mine, half); four<20>(a, b, c, d, mine, half);
    four<24>(a, b, c, d, mine, half); four<28>(a, b, c, d, mine, half); four<32>(a, b, c, d, mine, half);
    four<36>(a, b, c, d, mine, half); four<40>(a, b, c, d, mine, half); four<44>(a, b, c, d, mine, half);
    four<48>(a, b, c, d, mine, half); four<52>(a, b, c, d, mine, half); four<56>(a, b, c, d, mine, half);
    four<60>(a, b, c, d, mine, half);
    h[0] += a; h[1] += b; h[2] += c; h[3] += d;
  }
};

__global__ void __launch_bounds__(128) etag_kernel(const uint8_t* __restrict__ md5s,
                                                   const uint64_t* __restrict__ prefix, uint32_t n_objects,
                                                   uint8_t* __restrict__ out) {
  const uint32_t obj = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (obj >= n_objects) return;  // warp-uniform
  const uint64_t first = prefix[obj], parts = prefix[obj + 1] - first;
  const uint32_t* msg = reinterpret_cast<const uint32_t*>(md5s + first * 16);
  const uint64_t nwords = parts * 4;        // message words
  const uint64_t nblocks = parts / 4 + 1;   // data blocks + the one padding block
  const uint64_t bits = parts * 128;
  const uint64_t len_word = (nblocks - 1) * 16 + 14;
  uint32_t h[4];
  Md5::init(h);
  for (uint64_t wb = 0; wb < nblocks; wb += 2) {
    const uint64_t g = wb * 16 + lane;
    uint32_t mine = 0;
    if (g < nwords) mine = __ldg(msg + g);
    else if (g == nwords) mine = 0x80u;
    if (g == len_word) mine = (uint32_t)bits;
    if (g == len_word + 1) mine = (uint32_t)(bits >> 32);
    Md5Shfl::block(h, mine, 0);
    if (wb + 1 < nblocks) Md5Shfl::block(h, mine, 1);
  }
  const uint32_t v = lane == 0 ? h[0] : lane == 1 ? h[1] : lane == 2 ? h[2] : h[3];
  if (lane < 4) reinterpret_cast<uint32_t*>(out + (size_t)obj * 16)[lane] = v;
}

// ------------------------------------------------------------------------------------ synthetic data
__device__ __forceinline__ uint64_t splitmix64(uint64_t x) {
  x += 0x9e3779b97f4a7c15ull;
  x = (x ^ (x >> 30)) * 0xbf58476d1ce4e5b9ull;
  x = (x ^ (x >> 27)) * 0x94d049bb133111ebull;
  return x ^ (x >> 31);
}

__global__ void __launch_bounds__(256) synth_kernel(uint8_t* __restrict__ dst, uint64_t first_msg, uint64_t n,
                                                    uint64_t len, uint64_t stride, uint64_t seed) {
  const uint64_t words_per_msg = (len + 7) / 8;
  const uint64_t total = n * words_per_msg;
  for (uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total;
       t += (uint64_t)gridDim.x * blockDim.x) {
    const uint64_t k = t / words_per_msg, j = t - k * words_per_msg;
    const uint64_t w = splitmix64(seed + ((first_msg + k) << 40) + j);
    uint8_t* p = dst + k * stride + j * 8;
    const uint64_t left = len - j * 8;
    if (left >= 8 && (reinterpret_cast<uintptr_t>(p) & 7) == 0) {
      *reinterpret_cast<uint64_t*>(p) = w;
    } else {
      for (uint64_t b = 0; b < (left < 8 ? left : 8); b++) p[b] = (uint8_t)(w >> (8 * b));
    }
  }
}

// ------------------------------------------------------------------------------------ dispatch
// cudaFuncSetAttribute is per device: remember, per kernel instantiation, which devices have been set up.
template <class KernelPtr>
static cudaError_t allow_smem(KernelPtr kernel, bool* done /* [64] */, int bytes) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
  if (!done[dev]) {
    if (bytes > 48 * 1024) {
      e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
      if (e != cudaSuccess) return e;
    }
    // Every digest kernel asks for the same (maximum) shared-memory carve-out: an SM has to drain before its
    // L1 / shared split can change, which serialised long kernels of different submits that met on one SM.
    e = cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared);
    if (e != cudaSuccess) return e;
    done[dev] = true;
  }
  return cudaSuccess;
}

using launch_fn = cudaError_t (*)(const DigestArgs&, cudaStream_t);

template <class A, int CH, int NST>
cudaError_t launch_variant(const DigestArgs& a, cudaStream_t st) {
  constexpr int smem = Staging<CH, NST>::WARP_BYTES;  // <= 34 KiB for every compiled variant
  static bool done[64] = {};
  cudaError_t e0 = allow_smem(digest_kernel<A, CH, NST>, done, smem);
  if (e0 != cudaSuccess) return e0;
  const uint32_t grid = (a.n + 31u) / 32u;
  digest_kernel<A, CH, NST><<<grid, 32, smem, st>>>(a);
  return cudaGetLastError();
}

template <class A, int K, int NST>
cudaError_t launch_rows(const DigestArgs& a, cudaStream_t st) {
  constexpr int smem = Tiles<K, NST>::WARP_BYTES;
  static bool done[64] = {};
  cudaError_t e0 = allow_smem(digest_rows_kernel<A, K, NST>, done, smem);
  if (e0 != cudaSuccess) return e0;
  digest_rows_kernel<A, K, NST><<<(a.n + 31u) / 32u, 32, smem, st>>>(a);
  return cudaGetLastError();
}

struct RowsVariant { int k, nst; launch_fn fn[3]; };
#define TH_ROWS(K, N) {K, N, {launch_rows<Sha1, K, N>, launch_rows<Md5, K, N>, launch_rows<Sha256, K, N>}}
static const RowsVariant g_rows_variants[] = {TH_ROWS(1, 3), TH_ROWS(1, 4), TH_ROWS(2, 2), TH_ROWS(2, 3), TH_ROWS(4, 2)};
constexpr int kNumRowsVariants = (int)(sizeof(g_rows_variants) / sizeof(g_rows_variants[0]));

// ---- tensor-map path (equal-length, fixed-stride batches) -------------------------------------------
using EncodeTiledFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_tiled() {
  static EncodeTiledFn fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      p = nullptr;
    return reinterpret_cast<EncodeTiledFn>(p);
  }();
  return fn;
}

// Byte tensor [n rows][len bytes], row pitch `stride`; box {128 B, 32 rows}, 128-byte swizzle.
static bool make_tmap(CUtensorMap* m, const void* base, uint64_t len, uint64_t stride, uint32_t n) {
  EncodeTiledFn enc = encode_tiled();
  if (!enc) return false;
  const cuuint64_t gdim[2] = {len, n};
  const cuuint64_t gstr[1] = {n > 1 ? stride : ((len + 15) & ~15ull)};
  const cuuint32_t box[2] = {128, 32};
  const cuuint32_t estr[2] = {1, 1};
  return enc(m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void*>(base), gdim, gstr, box, estr,
             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

struct TileOpts {
  int path = 0;        // 0 auto, 1 per-lane bulk copies (any layout), 2 tile kernel, 3 sliced kernel, 4 assist kernel
  int assist_variant = 0;       // tuning: which (tiles, stages, ring slots) instantiation path 4 uses
  int assist_max_groups = 296;  // auto: chain + helper warp per group while two sub-partitions per group exist
  int workers = 12;    // sliced kernel: worker warps per SM (multiple of 4)
  int rounds = 32;     // sliced kernel: target scheduling rounds (units per worker)
  int slice_min_groups = 12;  // sliced kernel: cut chains into segments when a CTA has more groups than this
  int sliced_min_groups = 592;  // auto: use the sliced kernel from this many 32-message groups
  int tile_many_groups = 8192;  // auto: SHA-1 / SHA-256 batches with at least this many groups go back to one-warp CTAs
                                // (many waves balance themselves and measured 1-12 % faster than the persistent kernel)
  int md5_lone = 1;             // MD5 with at most one group per SM sub-partition: the run-ahead stream (tile_stream_lone)
  int parts_variant = 0;        // digest_parts_kernel tile geometry: 0 = 2 tiles x 2 stages, 1 = 4 x 2, 2 = 2 x 3 (tuning)
  int fuse_md5_sha256 = 1;      // th_md5_sha256_batch: 1 = one launch (digest_parts_kernel) where the SHA-256 plan is "assist"
  int ragged_rows = 1;          // per-lane kernels: 1 = cooperative cp.async row tiles, 0 = one bulk copy per lane
  int rows_variant[3] = {2, 2, 2};  // index into g_rows_variants per digest
};
static TileOpts g_opts;

// Launches with at most one group per SM ask for this much shared memory although they need less: then no more
// than two such CTAs (of any concurrently running digest kernel) fit on an SM, so their at most four long-running
// warps get a sub-partition each.  Without it the block scheduler stacked chains of different submits on the same
// sub-partitions (SHA-256 kernels of 1.05 s stretched to 2.1 and 3.1 s) while most of the chip was idle.
constexpr int kRoomySmem = 112 * 1024;

using tile_fn = cudaError_t (*)(const CUtensorMap&, const TileArgs&, uint32_t grid, uint32_t block, size_t smem, cudaStream_t);

template <class A, int K, int NST, bool LONE = false>
cudaError_t launch_tile(const CUtensorMap& m, const TileArgs& a, uint32_t grid, uint32_t, size_t pad, cudaStream_t st) {
  constexpr int smem = Tiles<K, NST>::WARP_BYTES + 8 * NST;
  static bool done[64] = {};
  cudaError_t e0 = allow_smem(digest_tile_kernel<A, K, NST, LONE>, done, smem > kRoomySmem ? smem : kRoomySmem);
  if (e0 != cudaSuccess) return e0;
  digest_tile_kernel<A, K, NST, LONE><<<grid, 32, pad > (size_t)smem ? pad : (size_t)smem, st>>>(m, a);
  return cudaGetLastError();
}

template <class A, int K, int NST>
cudaError_t launch_sliced(const CUtensorMap& m, const TileArgs& a, uint32_t grid, uint32_t block, size_t smem, cudaStream_t st) {
  static bool done[64] = {};
  cudaError_t e0 = allow_smem(digest_sliced_kernel<A, K, NST>, done, 227 * 1024);
  if (e0 != cudaSuccess) return e0;
  digest_sliced_kernel<A, K, NST><<<grid, block, smem, st>>>(m, a);
  return cudaGetLastError();
}

template <class A, int K, int NST, int SS, int BPS = 1>
cudaError_t launch_assist(const CUtensorMap& m, const TileArgs& a, uint32_t grid, uint32_t, size_t pad, cudaStream_t st) {
  constexpr int smem = Tiles<K, NST>::WARP_BYTES + SS * BPS * Assist<A>::WPB * 128 + 8 * NST;
  static bool done[64] = {};
  cudaError_t e0 = allow_smem(digest_assist_kernel<A, K, NST, SS, BPS>, done, smem > kRoomySmem ? smem : kRoomySmem);
  if (e0 != cudaSuccess) return e0;
  digest_assist_kernel<A, K, NST, SS, BPS><<<grid, 64, pad > (size_t)smem ? pad : (size_t)smem, st>>>(m, a);
  return cudaGetLastError();
}

template <int K, int NST, int SS, int BPS, int GP>
constexpr int parts_smem() { return GP * (Tiles<K, NST>::WARP_BYTES + SS * BPS * Assist<Sha256>::WPB * 128) + GP * 64; }

template <int K, int NST, int SS, int BPS, int GP, int TRIPS>
cudaError_t launch_parts(const CUtensorMap& m, const TileArgs& a, uint32_t grid, size_t pad, cudaStream_t st) {
  constexpr int smem = parts_smem<K, NST, SS, BPS, GP>();
  static_assert(smem <= 227 * 1024, "digest_parts_kernel: shared memory");
  static_assert(GP * 2 * SS <= 15, "named barriers 1..15");
  static bool done[64] = {};
  cudaError_t e0 = allow_smem(digest_parts_kernel<K, NST, SS, BPS, GP, TRIPS>, done, smem > kRoomySmem ? smem : kRoomySmem);
  if (e0 != cudaSuccess) return e0;
  digest_parts_kernel<K, NST, SS, BPS, GP, TRIPS><<<grid, GP == 1 ? 96 : 256, pad > (size_t)smem ? pad : (size_t)smem, st>>>(m, a);
  return cudaGetLastError();
}

struct TileVariant { int k, nst; tile_fn tile[3]; tile_fn sliced[3]; };
#define TH_TILE_VARIANT(K, NST)                                                                         \
  { K, NST, { launch_tile<Sha1, K, NST>, launch_tile<Md5, K, NST>, launch_tile<Sha256, K, NST> },          \
    { launch_sliced<Sha1, K, NST>, launch_sliced<Md5, K, NST>, launch_sliced<Sha256, K, NST> } }
static const TileVariant g_tile_variants[] = {
    TH_TILE_VARIANT(1, 2), TH_TILE_VARIANT(1, 3), TH_TILE_VARIANT(1, 4),
    TH_TILE_VARIANT(2, 2), TH_TILE_VARIANT(2, 3), TH_TILE_VARIANT(4, 2),
};
constexpr int kNumTileVariants = sizeof(g_tile_variants) / sizeof(g_tile_variants[0]);
static int g_tile_selected[3] = {5, 5, 5};    // one-warp CTAs: 4 tiles x 2 stages
static int g_sliced_selected[3] = {3, 3, 3};  // persistent kernel: 2 tiles x 2 stages

// Chaining-value slots and hand-over counters are only needed when chains are cut into segments.
static size_t sliced_smem(int K, int NST, uint32_t W, uint32_t Gmax, int ns, uint32_t S) {
  const size_t base = (size_t)W * K * NST * 4096 + ((W * 8 * NST + 127) & ~127u);
  return S > 1 ? base + ((Gmax * 4 + 127) & ~127u) + (size_t)Gmax * ns * 128 : base;
}

struct Plan {
  int path = 1;  // 1 bulk, 2 tile, 3 sliced
  int k = 0, nst = 0;
  uint32_t grid = 0, block = 32, workers = 0, segc = 0, nseg = 1, groups_per_cta = 0;
  int variant = -1;  // index into g_tile_variants
  size_t smem = 0;
  bool lone_md5 = false;  // path 2, MD5, at most one group per sub-partition: digest_tile_kernel<Md5, 4, 3, LONE>
};

// Chooses the kernel for an equal-length, fixed-stride batch (path 1 = not eligible for tensor maps).
static Plan make_plan(int ai, uint64_t stride, uint64_t len, uint32_t n) {
  Plan pl;
  if (g_opts.path == 1) return pl;
  if (len < 1024 || len >= (1ull << 31) || (n > 1 && (stride & 15)) || (n > 1 && stride < len) || stride >= (1ull << 40)) return pl;
  if (!encode_tiled()) return pl;
  const int ns = ai == 0 ? 5 : ai == 1 ? 4 : 8;
  const uint32_t groups = (n + 31u) / 32u;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  // SHA-256 gains from one-warp CTAs from 8192 groups on (+4..12 %), SHA-1 only from twice that (+1 %)
  const bool many = ai != 1 && groups >= (uint32_t)g_opts.tile_many_groups * (ai == 0 ? 2u : 1u);
  const bool sliced = g_opts.path == 3 || (g_opts.path == 0 && groups >= (uint32_t)g_opts.sliced_min_groups && !many);
  if (sliced) {
    const TileVariant& v = g_tile_variants[g_sliced_selected[ai]];
    const uint32_t W = (uint32_t)g_opts.workers;
    const uint32_t grid = groups < (uint32_t)sms ? groups : (uint32_t)sms;
    const uint32_t Gmax = (groups + grid - 1) / grid;
    const uint32_t nch = (uint32_t)(len / (128u * v.k)) + 1u;
    uint32_t S = 1;
    if (Gmax > (uint32_t)g_opts.slice_min_groups && Gmax < 4 * W) {  // few chains per worker: cut them so that ~rounds units run per worker
      S = ((uint32_t)g_opts.rounds * W + Gmax - 1) / Gmax;
      const uint32_t smax = nch / 4 ? nch / 4 : 1;  // at least 4 chunks per segment
      if (S > smax) S = smax;
      if (S < 1) S = 1;
    }
    const size_t smem = sliced_smem(v.k, v.nst, W, Gmax, ns, S);
    if (smem <= 227 * 1024) {
      pl.path = 3; pl.k = v.k; pl.nst = v.nst; pl.grid = grid; pl.block = W * 32; pl.workers = W;
      pl.variant = g_sliced_selected[ai];
      pl.segc = (nch + S - 1) / S; pl.nseg = (nch + pl.segc - 1) / pl.segc; pl.groups_per_cta = Gmax; pl.smem = smem;
      return pl;
    }
    // chaining-value slots do not fit beside the tiles: use the one-warp-per-CTA kernel
  }
  if (ai != 1 && (g_opts.path == 4 || (g_opts.path == 0 && groups <= (uint32_t)g_opts.assist_max_groups))) {
    pl.path = 4; pl.k = 2; pl.nst = 2; pl.grid = groups; pl.block = 64;
    pl.smem = (size_t)2 * 2 * 4096 + 4 * (ai == 0 ? 80 : 64) * 128 + 16;
    return pl;
  }
  const TileVariant& v = g_tile_variants[g_opts.path == 0 && many ? g_sliced_selected[ai] : g_tile_selected[ai]];
  pl.path = 2; pl.k = v.k; pl.nst = v.nst; pl.grid = groups; pl.block = 32; pl.variant = int(&v - g_tile_variants);
  pl.smem = (size_t)v.k * v.nst * 4096 + 8 * v.nst;
  if (ai == 1 && g_opts.md5_lone && groups <= 4u * (uint32_t)sms) {  // MD5 chains alone on their sub-partitions
    pl.lone_md5 = true; pl.k = 4; pl.nst = 3; pl.smem = (size_t)4 * 3 * 4096 + 8 * 3;
  }
  return pl;
}

// Returns TH_OK if launched, 1 if this batch is not eligible (caller falls back to the bulk-copy kernel).
// d_out_md5 != nullptr (ai == 2 only): MD5 digests as well, from the same launch (digest_parts_kernel) — only where the
// SHA-256 plan is the chain + helper one; returns 1 for any other plan.
static int launch_uniform(int ai, const void* d_base, uint64_t stride, uint64_t len, uint32_t n, uint8_t* d_out,
                          const uint8_t* d_expected, uint32_t* d_ok_bits, cudaStream_t st, uint8_t* d_out_md5 = nullptr) {
  const Plan pl = make_plan(ai, stride, len, n);
  if (pl.path == 1) return 1;
  if (d_out_md5 && (pl.path != 4 || ai != 2)) return 1;
  CUtensorMap tmap;
  if (!make_tmap(&tmap, d_base, len, stride, n)) return 1;
  TileArgs a{len, n, pl.segc, pl.nseg, d_out, d_out_md5, d_expected, d_ok_bits, 1u, 0u};
  uint32_t grid = pl.grid;
  size_t pad = 0;
  if (pl.path != 3) {  // few CTAs: pad to one per SM and rotate by a per-launch salt (see spread_group)
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (grid <= (uint32_t)sms) {
      static std::atomic<uint32_t> launches{0};
      a.salt = (launches.fetch_add(1) * 37u) % (uint32_t)sms;
      grid = (uint32_t)sms;
      pad = kRoomySmem;
    }
  }
  if (pl.path == 4 && d_out_md5) {  // MD5 + SHA-256 in one launch (digest_parts_kernel)
    const uint32_t groups = (n + 31u) / 32u;
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaError_t e;
    if (groups <= (uint32_t)sms) {  // one group per CTA; `grid` is already padded to one CTA per SM and salted
      e = g_opts.parts_variant == 1   ? launch_parts<4, 2, 2, 2, 1, 4>(tmap, a, grid, pad, st)
          : g_opts.parts_variant == 2 ? launch_parts<2, 3, 2, 2, 1, 4>(tmap, a, grid, pad, st)
                                      : launch_parts<2, 2, 2, 2, 1, 4>(tmap, a, grid, pad, st);
    } else {                        // two groups per CTA, at most one CTA per group pair and SM
      a.salt = 0;
      e = g_opts.parts_variant == 1   ? launch_parts<4, 2, 2, 2, 2, 4>(tmap, a, (groups + 1u) / 2u, 0, st)
          : g_opts.parts_variant == 2 ? launch_parts<2, 3, 2, 2, 2, 4>(tmap, a, (groups + 1u) / 2u, 0, st)
                                      : launch_parts<2, 2, 2, 2, 2, 4>(tmap, a, (groups + 1u) / 2u, 0, st);
    }
    return e == cudaSuccess ? TH_OK : TH_ECUDA;
  }
  if (pl.path == 4) {
    cudaError_t e;
    switch (g_opts.assist_variant) {
      case 1: e = ai == 0 ? launch_assist<Sha1, 4, 2, 4>(tmap, a, grid, 64, pad, st) : launch_assist<Sha256, 4, 2, 4>(tmap, a, grid, 64, pad, st); break;
      case 2: e = ai == 0 ? launch_assist<Sha1, 4, 4, 4>(tmap, a, grid, 64, pad, st) : launch_assist<Sha256, 4, 4, 4>(tmap, a, grid, 64, pad, st); break;
      case 3: e = ai == 0 ? launch_assist<Sha1, 4, 4, 7>(tmap, a, grid, 64, pad, st) : launch_assist<Sha256, 4, 4, 7>(tmap, a, grid, 64, pad, st); break;
      case 4: e = ai == 0 ? launch_assist<Sha1, 2, 4, 2>(tmap, a, grid, 64, pad, st) : launch_assist<Sha256, 2, 4, 2>(tmap, a, grid, 64, pad, st); break;
      case 5: e = ai == 0 ? launch_assist<Sha1, 2, 2, 3, 2>(tmap, a, grid, 64, pad, st) : launch_assist<Sha256, 2, 2, 3, 2>(tmap, a, grid, 64, pad, st); break;  // 3 slots of 2 blocks
      case 7: e = ai == 0 ? launch_assist<Sha1, 2, 2, 4>(tmap, a, grid, 64, pad, st) : launch_assist<Sha256, 2, 2, 4>(tmap, a, grid, 64, pad, st); break;  // round 1's: 4 slots of 1 block
      default: e = ai == 0 ? launch_assist<Sha1, 2, 2, 2, 2>(tmap, a, grid, 64, pad, st) : launch_assist<Sha256, 2, 2, 2, 2>(tmap, a, grid, 64, pad, st); break;  // 2 slots of 2 blocks
    }
    return e == cudaSuccess ? TH_OK : TH_ECUDA;
  }
  if (pl.lone_md5) {
    const cudaError_t e = launch_tile<Md5, 4, 3, true>(tmap, a, grid, 32, pad, st);
    return e == cudaSuccess ? TH_OK : TH_ECUDA;
  }
  const TileVariant& v = g_tile_variants[pl.variant];
  cudaError_t e = pl.path == 3 ? v.sliced[ai](tmap, a, pl.grid, pl.block, pl.smem, st)
                               : v.tile[ai](tmap, a, grid, 32, pad, st);
  return e == cudaSuccess ? TH_OK : TH_ECUDA;
}

struct Variant { int ch, nst; launch_fn fn[3]; };
#define TH_VARIANT(CH, NST) \
  { CH, NST, { launch_variant<Sha1, CH, NST>, launch_variant<Md5, CH, NST>, launch_variant<Sha256, CH, NST> } }
static const Variant g_variants[] = {
    TH_VARIANT(64, 4),  TH_VARIANT(128, 2), TH_VARIANT(128, 3), TH_VARIANT(128, 4),
    TH_VARIANT(256, 2), TH_VARIANT(256, 3), TH_VARIANT(512, 2),
};
constexpr int kNumVariants = sizeof(g_variants) / sizeof(g_variants[0]);
static int g_selected[3] = {6, 2, 6};  // index into g_variants per algorithm: sha1, md5, sha256

static int alg_index(uint32_t alg) {
  return alg == TH_ALG_SHA1 ? 0 : alg == TH_ALG_MD5 ? 1 : alg == TH_ALG_SHA256 ? 2 : -1;
}

static int launch_digest(uint32_t alg, const void* d_base, const uint64_t* d_offsets, const uint64_t* d_lengths,
                         const uint32_t* d_order, uint64_t stride, uint64_t len, uint32_t n, uint8_t* d_out,
                         const uint8_t* d_expected, uint32_t* d_ok_bits, void* stream) {
  const int ai = alg_index(alg);
  if (ai < 0 || !d_out) return TH_EINVAL;
  if (n == 0) return TH_OK;
  if (!d_base && !(d_lengths == nullptr && len == 0)) return TH_EINVAL;
  if ((reinterpret_cast<uintptr_t>(d_base) & 15) != 0) return TH_EALIGN;
  if (!d_offsets && (stride & 15) != 0 && n > 1) return TH_EALIGN;
  if (!d_lengths && len >= (1ull << 37)) return TH_ETOOBIG;
  if (!d_offsets && !d_lengths && !d_order) {  // equal-length, fixed-stride batch -> tensor-map kernels
    const int rc = launch_uniform(ai, d_base, stride, len, n, d_out, d_expected, d_ok_bits, static_cast<cudaStream_t>(stream));
    if (rc <= 0) return rc;
  }
  DigestArgs a{static_cast<const uint8_t*>(d_base), d_offsets, d_lengths, d_order, stride, len, n, d_out,
               d_expected, d_ok_bits, 1u};
  cudaError_t e = g_opts.ragged_rows ? g_rows_variants[g_opts.rows_variant[ai]].fn[ai](a, static_cast<cudaStream_t>(stream))
                                    : g_variants[g_selected[ai]].fn[ai](a, static_cast<cudaStream_t>(stream));
  return e == cudaSuccess ? TH_OK : TH_ECUDA;
}

// ------------------------------------------------------------------- allocation-free ragged batches
// Scratch for th_digest_ragged_static: lives in the library image, one copy per GPU that loads it,
// until the process exits.  Host code may not treat `g_static_scratch` as a device address — it is
// only a handle here; the real address comes from cudaGetSymbolAddress (static_scratch_ptr below).
constexpr uint32_t kLenClasses = 128;  // 4 classes per power of two of the block count
struct StaticScratch {
  uint32_t count[kLenClasses];   // messages per length class      (zeroed at the start of every call)
  uint32_t cursor[kLenClasses];  // next free place in each class   (zeroed at the start of every call)
  uint32_t order[TH_STATIC_MAX_MSGS];  // thread slot -> message index, longest classes first
};
static __device__ __align__(256) unsigned char g_static_scratch[sizeof(StaticScratch)];

__device__ __forceinline__ uint32_t len_class(uint64_t len) {
  const uint64_t blocks = (len >> 6) + 1;                     // >= 1
  const int e = 63 - __clzll((long long)blocks);              // floor(log2(blocks)), 0..58
  const uint32_t frac = e >= 2 ? (uint32_t)((blocks >> (e - 2)) & 3u) : (uint32_t)((blocks << (2 - e)) & 3u);
  const uint32_t c = (uint32_t)e * 4u + frac;
  return c < kLenClasses ? c : kLenClasses - 1;
}

__global__ void __launch_bounds__(256) class_count_kernel(const uint64_t* __restrict__ lengths, uint32_t n,
                                                          StaticScratch* s) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) atomicAdd(&s->count[len_class(lengths[i])], 1u);
}

__global__ void __launch_bounds__(256) class_place_kernel(const uint64_t* __restrict__ lengths, uint32_t n,
                                                          StaticScratch* s) {
  __shared__ uint32_t base[kLenClasses];
  if (threadIdx.x == 0) {  // longest class first
    uint32_t acc = 0;
    for (int c = (int)kLenClasses - 1; c >= 0; c--) { base[c] = acc; acc += s->count[c]; }
  }
  __syncthreads();
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    const uint32_t c = len_class(lengths[i]);
    s->order[base[c] + atomicAdd(&s->cursor[c], 1u)] = i;
  }
}

// Device address of the scratch array for the current device, looked up once per device.
static cudaError_t static_scratch_ptr(StaticScratch** out) {
  static StaticScratch* cached[64] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
  if (!cached[dev]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_static_scratch);
    if (e != cudaSuccess) return e;
    cached[dev] = static_cast<StaticScratch*>(p);
  }
  *out = cached[dev];
  return cudaSuccess;
}

}  // namespace th

// ----------------------------------------------------------------------------------------- C ABI
extern "C" {

int th_sha1_batch(const void* d_base, const uint64_t* d_offsets, const uint64_t* d_lengths,
                  const uint32_t* d_order, uint64_t stride, uint64_t len, uint32_t n, uint8_t* d_out20,
                  const uint8_t* d_expected20, uint32_t* d_ok_bits, void* stream) {
  return th::launch_digest(TH_ALG_SHA1, d_base, d_offsets, d_lengths, d_order, stride, len, n, d_out20,
                           d_expected20, d_ok_bits, stream);
}

int th_md5_batch(const void* d_base, const uint64_t* d_offsets, const uint64_t* d_lengths,
                 const uint32_t* d_order, uint64_t stride, uint64_t len, uint32_t n, uint8_t* d_out16,
                 void* stream) {
  return th::launch_digest(TH_ALG_MD5, d_base, d_offsets, d_lengths, d_order, stride, len, n, d_out16, nullptr,
                           nullptr, stream);
}

int th_sha256_batch(const void* d_base, const uint64_t* d_offsets, const uint64_t* d_lengths,
                    const uint32_t* d_order, uint64_t stride, uint64_t len, uint32_t n, uint8_t* d_out32,
                    void* stream) {
  return th::launch_digest(TH_ALG_SHA256, d_base, d_offsets, d_lengths, d_order, stride, len, n, d_out32,
                           nullptr, nullptr, stream);
}

int th_md5_sha256_fused(uint64_t stride, uint64_t len, uint32_t n) {
  if (!th::g_opts.fuse_md5_sha256 || n == 0) return 0;
  return th::make_plan(2, stride, len, n).path == 4 ? 1 : 0;
}

int th_md5_sha256_batch(const void* d_base, uint64_t stride, uint64_t len, uint32_t n, uint8_t* d_out16,
                        uint8_t* d_out32, void* stream) {
  if (!d_out16 || !d_out32) return TH_EINVAL;
  if (n == 0) return TH_OK;
  if (!d_base) return TH_EINVAL;
  if ((reinterpret_cast<uintptr_t>(d_base) & 15) != 0 || ((stride & 15) != 0 && n > 1)) return TH_EALIGN;
  if (len >= (1ull << 37)) return TH_ETOOBIG;
  if (th_md5_sha256_fused(stride, len, n)) {
    const int rc = th::launch_uniform(2, d_base, stride, len, n, d_out32, nullptr, nullptr, static_cast<cudaStream_t>(stream), d_out16);
    if (rc <= 0) return rc;
  }
  const int rc = th_md5_batch(d_base, nullptr, nullptr, nullptr, stride, len, n, d_out16, stream);
  return rc != TH_OK ? rc : th_sha256_batch(d_base, nullptr, nullptr, nullptr, stride, len, n, d_out32, stream);
}

int th_etag_combine_device(const uint8_t* d_md5s, const uint64_t* d_part_prefix, uint32_t n_objects,
                           uint8_t* d_out16, void* stream) {
  if (!d_part_prefix || !d_out16) return TH_EINVAL;
  if (n_objects == 0) return TH_OK;
  if ((reinterpret_cast<uintptr_t>(d_md5s) & 3) != 0) return TH_EALIGN;
  const uint32_t warps = 4;
  th::etag_kernel<<<(n_objects + warps - 1) / warps, warps * 32, 0, static_cast<cudaStream_t>(stream)>>>(
      d_md5s, d_part_prefix, n_objects, d_out16);
  return cudaGetLastError() == cudaSuccess ? TH_OK : TH_ECUDA;
}

int th_synth_fill(void* d_dst, uint64_t first_msg, uint64_t n, uint64_t len, uint64_t stride, uint64_t seed,
                  void* stream) {
  if (!d_dst) return TH_EINVAL;
  if (n == 0 || len == 0) return TH_OK;
  const uint64_t total = n * ((len + 7) / 8);
  uint64_t blocks = (total + 255) / 256;
  if (blocks > 148ull * 64) blocks = 148ull * 64;
  th::synth_kernel<<<(unsigned)blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<uint8_t*>(d_dst), first_msg, n, len, stride, seed);
  return cudaGetLastError() == cudaSuccess ? TH_OK : TH_ECUDA;
}

int th_digest_ragged_static(uint32_t alg, const void* d_base, const uint64_t* d_offsets, const uint64_t* d_lengths,
                            uint32_t n, uint8_t* d_out, const uint8_t* d_expected20, uint32_t* d_ok_bits,
                            void* stream) {
  if (th::alg_index(alg) < 0 || !d_out || !d_offsets || !d_lengths) return TH_EINVAL;
  if (n > TH_STATIC_MAX_MSGS) return TH_ETOOBIG;  // the built-in scratch holds TH_STATIC_MAX_MSGS slots
  if (n == 0) return TH_OK;
  // everything the host can check is checked before the first enqueue, so a refused call leaves the stream untouched
  if (!d_base) return TH_EINVAL;
  if ((reinterpret_cast<uintptr_t>(d_base) & 15) != 0) return TH_EALIGN;
  th::StaticScratch* s = nullptr;
  // First call on this device: the symbol lookup (like any first launch) makes the driver load this library's
  // device code and place the scratch array in device memory.  That is the one moment the call can run out of memory.
  const cudaError_t e0 = th::static_scratch_ptr(&s);
  if (e0 == cudaErrorMemoryAllocation) { cudaGetLastError(); return TH_ENOMEM; }
  if (e0 != cudaSuccess) return TH_ECUDA;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  // the array still holds the previous call's counters: they must start from zero
  if (cudaMemsetAsync(s, 0, 2 * th::kLenClasses * sizeof(uint32_t), st) != cudaSuccess) return TH_ECUDA;
  const bool verify = alg == TH_ALG_SHA1 && d_expected20 && d_ok_bits;
  if (verify && cudaMemsetAsync(d_ok_bits, 0, ((n + 31u) / 32u) * 4u, st) != cudaSuccess) return TH_ECUDA;
  const uint32_t blocks = (n + 255u) / 256u;
  th::class_count_kernel<<<blocks, 256, 0, st>>>(d_lengths, n, s);
  th::class_place_kernel<<<blocks, 256, 0, st>>>(d_lengths, n, s);
  const cudaError_t e1 = cudaGetLastError();
  if (e1 == cudaErrorMemoryAllocation) return TH_ENOMEM;
  if (e1 != cudaSuccess) return TH_ECUDA;
  return th::launch_digest(alg, d_base, d_offsets, d_lengths, s->order, 0, 0, n, d_out,
                           verify ? d_expected20 : nullptr, verify ? d_ok_bits : nullptr, stream);
}

int th_set_staging(uint32_t alg, int chunk_bytes, int stages) {
  int rc = TH_EINVAL;
  for (int v = 0; v < th::kNumVariants; v++) {
    if (th::g_variants[v].ch == chunk_bytes && th::g_variants[v].nst == stages) {
      for (int a = 0; a < 3; a++)
        if (alg & (1u << a)) { th::g_selected[a] = v; rc = TH_OK; }
    }
  }
  return rc;
}

int th_set_option(const char* name, long long value) {
  if (!name) return TH_EINVAL;
  const std::string k(name);
  if (k == "path") { if (value < 0 || value > 4) return TH_EINVAL; th::g_opts.path = (int)value; return TH_OK; }
  if (k == "assist_variant") { if (value < 0 || value > 7 || value == 6) return TH_EINVAL; th::g_opts.assist_variant = (int)value; return TH_OK; }
  if (k == "assist_max_groups") { if (value < 0) return TH_EINVAL; th::g_opts.assist_max_groups = (int)value; return TH_OK; }
  if (k == "workers") { if (value < 4 || value > 16 || value % 4) return TH_EINVAL; th::g_opts.workers = (int)value; return TH_OK; }
  if (k == "slice_min_groups") { if (value < 0) return TH_EINVAL; th::g_opts.slice_min_groups = (int)value; return TH_OK; }
  if (k == "rounds") { if (value < 1 || value > 4096) return TH_EINVAL; th::g_opts.rounds = (int)value; return TH_OK; }
  if (k == "tile_many_groups") { if (value < 1) return TH_EINVAL; th::g_opts.tile_many_groups = (int)value; return TH_OK; }
  if (k == "sliced_min_groups") { if (value < 1) return TH_EINVAL; th::g_opts.sliced_min_groups = (int)value; return TH_OK; }
  if (k == "ragged_rows") { if (value < 0 || value > 1) return TH_EINVAL; th::g_opts.ragged_rows = (int)value; return TH_OK; }
  if (k == "md5_lone") { if (value < 0 || value > 1) return TH_EINVAL; th::g_opts.md5_lone = (int)value; return TH_OK; }
  if (k == "parts_variant") { if (value < 0 || value > 2) return TH_EINVAL; th::g_opts.parts_variant = (int)value; return TH_OK; }
  if (k == "fuse_md5_sha256") { if (value < 0 || value > 1) return TH_EINVAL; th::g_opts.fuse_md5_sha256 = (int)value; return TH_OK; }
  if (k == "rows_variant") {
    if (value < 0 || value >= th::kNumRowsVariants) return TH_EINVAL;
    for (int a = 0; a < 3; a++) th::g_opts.rows_variant[a] = (int)value;
    return TH_OK;
  }
  if (k == "tile_k" || k == "tile_stages") return TH_EINVAL;  // use th_set_tiling
  return TH_EINVAL;
}

int th_set_tiling(uint32_t alg, int tiles_per_stage, int stages) {
  int rc = TH_EINVAL;
  for (int v = 0; v < th::kNumTileVariants; v++) {
    if (th::g_tile_variants[v].k == tiles_per_stage && th::g_tile_variants[v].nst == stages) {
      for (int a = 0; a < 3; a++)
        if (alg & (1u << a)) { th::g_tile_selected[a] = v; th::g_sliced_selected[a] = v; rc = TH_OK; }
    }
  }
  return rc;
}

int th_plan(uint32_t alg, uint64_t len, uint64_t stride, uint32_t n, char* out, size_t cap) {
  const int ai = th::alg_index(alg);
  if (ai < 0 || !out || cap == 0) return TH_EINVAL;
  const th::Plan pl = th::make_plan(ai, stride, len, n);
  if (pl.path == 1) {
    const th::Variant& v = th::g_variants[th::g_selected[ai]];
    const th::RowsVariant& r = th::g_rows_variants[th::g_opts.rows_variant[ai]];
    if (th::g_opts.ragged_rows)
      snprintf(out, cap, "rows: one-warp CTAs, cooperative cp.async into {128B x 32 msgs} x %d tiles x %d stages, grid %u", r.k, r.nst, (n + 31u) / 32u);
    else
      snprintf(out, cap, "bulk: one-warp CTAs, per-lane cp.async.bulk, %d B x %d stages, grid %u", v.ch, v.nst, (n + 31u) / 32u);
  } else if (pl.path == 4) {
    snprintf(out, cap,
             "assist: %u CTAs x (chain warp + schedule-helper warp), TMA 2-D {128B x 32 msgs} x %d tiles x %d stages, "
             "schedule ring of 2 slots x 2 blocks, smem %zu B", pl.grid, pl.k, pl.nst, pl.smem);
  } else if (pl.path == 2) {
    snprintf(out, cap, "tile%s: one-warp CTAs, TMA 2-D {128B x 32 msgs} x %d tiles x %d stages, grid %u, smem %zu B",
             pl.lone_md5 ? " (lone chains: loads of the next block run ahead of each compression)" : "", pl.k, pl.nst, pl.grid, pl.smem);
  } else {
    snprintf(out, cap,
             "sliced: persistent %u CTAs x %u warps, TMA 2-D {128B x 32 msgs} x %d tiles x %d stages, <=%u groups/CTA, "
             "%u segments of %u chunks, smem %zu B",
             pl.grid, pl.workers, pl.k, pl.nst, pl.groups_per_cta, pl.nseg, pl.segc, pl.smem);
  }
  return TH_OK;
}

int th_get_staging(uint32_t alg, int* chunk_bytes, int* stages) {
  const int ai = th::alg_index(alg);
  if (ai < 0) return TH_EINVAL;
  if (chunk_bytes) *chunk_bytes = th::g_variants[th::g_selected[ai]].ch;
  if (stages) *stages = th::g_variants[th::g_selected[ai]].nst;
  return TH_OK;
}

}  // extern "C"
