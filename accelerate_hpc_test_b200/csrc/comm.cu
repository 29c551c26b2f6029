// Peer-memory collectives over NVLink 5 / NVSwitch, issued from inside the kernels (no NCCL on the
// data path).  Every rank maps every peer's symmetric buffer (runtime.cu: CUDA IPC); `peers[p]` is
// the base address of rank p's buffer as seen from this GPU.  Ranks order their accesses with
// monotonically increasing epochs written into each other's signal pads with st.release.sys and
// polled with ld.acquire.sys.
//
// SURVEY.md §2.6 collectives C1 (DDP all-reduce, fused with cast+scale), C2 (FSDP2 all-gather), C3
// (FSDP2 reduce-scatter, fused here with the sharded AdamW update), C6 (Ulysses all-to-all).  In the
// reference these are ncclAllReduce / ncclAllGather / ncclReduceScatter / all_to_all_single issued by
// torch DDP / FSDP2 / DeepSpeed (reference src/accelerate/accelerator.py:1865, utils/fsdp_utils.py:689).
//
// Signal pad layout (start of every symmetric buffer): uint32 flags[NUM_CHANNELS][MAX_RANKS];
// flags[c][s] on rank d is written only by rank s.  All waits are bounded (device-side timeout ->
// trap with a diagnostic) so a missing peer turns into an error, never into a hung GPU.
#include "common.cuh"
#include <string.h>
#include <stdlib.h>

namespace b200 {

static constexpr int MAX_RANKS = 8;
static constexpr int NUM_CHANNELS = 2048;
__device__ unsigned long long g_wait_timeout_ns = 60ull * 1000ull * 1000ull * 1000ull;

struct Peers {
  uint8_t* base[MAX_RANKS];
};

__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

__device__ __forceinline__ uint32_t* flag_ptr(uint8_t* base, int channel, int src) {
  return reinterpret_cast<uint32_t*>(base) + channel * MAX_RANKS + src;
}

// tell every peer (and myself) that `rank` reached `epoch` on `channel`; one thread per destination
__device__ __forceinline__ void signal_all(const Peers& peers, int rank, int world, int channel, uint32_t epoch) {
  if (threadIdx.x < world) st_release_sys(flag_ptr(peers.base[threadIdx.x], channel, rank), epoch);
}

// wait until every rank has reached `epoch` on `channel` (reads only my own pad); whole CTA syncs after
__device__ __forceinline__ void wait_all(const Peers& peers, int rank, int world, int channel, uint32_t epoch) {
  if (threadIdx.x < world) {
    const uint32_t* f = flag_ptr(peers.base[rank], channel, threadIdx.x);
    const unsigned long long t0 = global_ns();
    while ((int32_t)(ld_acquire_sys(f) - epoch) < 0) {
      __nanosleep(64);
      if (global_ns() - t0 > g_wait_timeout_ns) {
        printf("b200 comm: rank %d timed out waiting for rank %d on channel %d (epoch %u, saw %u)\n", rank,
               (int)threadIdx.x, channel, epoch, ld_relaxed_sys(f));
        __trap();
      }
    }
  }
  __syncthreads();
}

// last-CTA election on a local counter; returns true in exactly one CTA, after all CTAs arrived
__device__ __forceinline__ bool last_cta(uint32_t* counter) {
  __shared__ uint32_t is_last;
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const uint32_t prev = atomicAdd(counter, 1u);
    is_last = (prev == gridDim.x - 1) ? 1u : 0u;
    if (is_last) *counter = 0;  // re-arm for the next launch
  }
  __syncthreads();
  return is_last != 0;
}

__device__ __forceinline__ uint4 ld_peer_v4(const void* p) {
  uint4 r;
  asm volatile("ld.global.relaxed.sys.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p)
               : "memory");
  return r;
}

__device__ __forceinline__ void accum_bf16x8(float (&acc)[8], const uint4& v) {
  float2 t;
  t = bf2_to_f2(v.x); acc[0] += t.x; acc[1] += t.y;
  t = bf2_to_f2(v.y); acc[2] += t.x; acc[3] += t.y;
  t = bf2_to_f2(v.z); acc[4] += t.x; acc[5] += t.y;
  t = bf2_to_f2(v.w); acc[6] += t.x; acc[7] += t.y;
}

// ------------------------------------------------------------------------------ barrier
__global__ void barrier_kernel(Peers peers, int rank, int world, int channel, uint32_t epoch) {
  signal_all(peers, rank, world, channel, epoch);
  wait_all(peers, rank, world, channel, epoch);
}

// ------------------------------------------------------------------------------ all-gather (pull)
// out[p * shard_bytes ...] = peer p's bytes at [shard_off, shard_off + shard_bytes).
// LSU loads over NVLink top out at a few GB/s per SM (outstanding-request limit), so the copy is
// driven by the TMA engine instead: one thread per CTA streams 16 KB chunks peer -> smem ring ->
// local memory with cp.async.bulk, several chunks in flight.  ~16-32 CTAs saturate the link and the
// SMs stay free for the GEMMs running next to it.
// If wait_channel >= 0 peer data is read only once flags[wait_channel][p] >= wait_epoch.
// The ring is kept under ~25 KB so a copy CTA co-resides with a persistent GEMM CTA (198 KB of smem)
// on every SM: the gather then runs *beside* the tensor-core work instead of waiting for an SM.
static constexpr int AG_CHUNK = 8192;
static constexpr int AG_STAGES = 3;
static constexpr int AG_AHEAD = 2;
static constexpr int AG_SMEM = AG_CHUNK * AG_STAGES + 128 + 128;

__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(smem_dst)),
               "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void bulk_s2g(void* gdst, const void* smem_src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(smem_src)),
               "r"(bytes)
               : "memory");
}

// Segment table of a sharded unit.  A unit's shard is the concatenation of one piece per segment; in the gathered ("full")
// buffer segment i occupies [full_off, full_off + world * bytes) with rank q's piece at full_off + q * bytes.  A 2-D weight
// sharded along dim 0 is one segment (piece = a block of rows), so the gathered buffer holds it in its natural layout.
static constexpr int MAX_SEGS = 32;
struct Segs {
  long long shard_off[MAX_SEGS];  // byte offset of the piece inside the shard
  long long full_off[MAX_SEGS];   // byte offset of the segment inside the gathered buffer
  long long bytes[MAX_SEGS];      // bytes per piece
  int n;
};

__global__ void __launch_bounds__(32)
allgather_pull_kernel(Peers peers, int rank, int world, long long shard_off, Segs segs,
                      uint8_t* __restrict__ out, int wait_channel, uint32_t wait_epoch) {
  extern __shared__ uint8_t ag_smem_raw[];
  uint8_t* ring = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(ag_smem_raw) + 127) & ~(uintptr_t)127);
  uint64_t* bars = reinterpret_cast<uint64_t*>(ring + AG_CHUNK * AG_STAGES);
  if (wait_channel >= 0) wait_all(peers, rank, world, wait_channel, wait_epoch);
  if (threadIdx.x == 0) {
    for (int i = 0; i < AG_STAGES; ++i) mbar_init(&bars[i], 1);
    fence_barrier_init();
  }
  __syncwarp();
  if (threadIdx.x != 0) return;
  long long chunks_per_shard = 0;
  for (int i = 0; i < segs.n; ++i) chunks_per_shard += (segs.bytes[i] + AG_CHUNK - 1) / AG_CHUNK;
  const long long total = chunks_per_shard * world;
  // chunk g of this CTA: g = blockIdx.x + k * gridDim.x ; peers visited starting at my right neighbour
  const long long n = (total > blockIdx.x) ? (total - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
  auto locate = [&](long long k, const uint8_t*& src, uint8_t*& dst, uint32_t& bytes) {
    const long long g = blockIdx.x + k * (long long)gridDim.x;
    const int slot = (int)(g / chunks_per_shard);
    long long c = g - (long long)slot * chunks_per_shard;
    const int p = (rank + 1 + slot) % world;
    int si = 0;
    for (; si + 1 < segs.n; ++si) {
      const long long sc = (segs.bytes[si] + AG_CHUNK - 1) / AG_CHUNK;
      if (c < sc) break;
      c -= sc;
    }
    const long long o = c * AG_CHUNK;
    const long long left = segs.bytes[si] - o;
    bytes = (uint32_t)(left < AG_CHUNK ? left : AG_CHUNK);
    src = peers.base[p] + shard_off + segs.shard_off[si] + o;
    dst = out + segs.full_off[si] + (long long)p * segs.bytes[si] + o;
  };
  auto issue_load = [&](long long k) {
    const uint8_t* src; uint8_t* dst; uint32_t bytes;
    locate(k, src, dst, bytes);
    const int st = (int)(k % AG_STAGES);
    mbar_arrive_expect_tx(&bars[st], bytes);
    bulk_g2s(ring + st * AG_CHUNK, src, bytes, &bars[st]);
  };
  for (long long k = 0; k < AG_AHEAD && k < n; ++k) issue_load(k);
  for (long long k = 0; k < n; ++k) {
    if (k + AG_AHEAD < n) {
      // the ring slot of chunk k+AHEAD was last read by the store of chunk k+AHEAD-STAGES
      asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(AG_STAGES - AG_AHEAD - 1) : "memory");
      issue_load(k + AG_AHEAD);
    }
    const int st = (int)(k % AG_STAGES);
    mbar_wait(&bars[st], (uint32_t)((k / AG_STAGES) & 1));
    const uint8_t* src; uint8_t* dst; uint32_t bytes;
    locate(k, src, dst, bytes);
    bulk_s2g(dst, ring + st * AG_CHUNK, bytes);
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
  }
  asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

// All-reduce phase 2: slice p of the buffer (already reduced by its owner p) is pulled from p into the same place of my
// buffer.  Same TMA ring as the all-gather; the last CTA closes with a cross-rank barrier so that nobody reuses its buffer
// while a peer is still copying out of it.
__global__ void __launch_bounds__(32)
allgather_inplace_kernel(Peers peers, int rank, int world, long long off, long long slice_bytes, int ch_done, uint32_t epoch,
                         uint32_t* __restrict__ counter) {
  extern __shared__ uint8_t ag_smem_raw[];
  uint8_t* ring = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(ag_smem_raw) + 127) & ~(uintptr_t)127);
  uint64_t* bars = reinterpret_cast<uint64_t*>(ring + AG_CHUNK * AG_STAGES);
  if (threadIdx.x == 0) {
    for (int i = 0; i < AG_STAGES; ++i) mbar_init(&bars[i], 1);
    fence_barrier_init();
  }
  __syncwarp();
  if (threadIdx.x == 0) {
    const long long cps = (slice_bytes + AG_CHUNK - 1) / AG_CHUNK;
    const long long total = cps * (world - 1);
    const long long n = (total > blockIdx.x) ? (total - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    auto locate = [&](long long k, const uint8_t*& src, uint8_t*& dst, uint32_t& bytes) {
      const long long g = blockIdx.x + k * (long long)gridDim.x;
      const int slot = (int)(g / cps);
      const long long c = g - (long long)slot * cps;
      const int p = (rank + 1 + slot) % world;
      const long long o = c * AG_CHUNK;
      const long long left = slice_bytes - o;
      bytes = (uint32_t)(left < AG_CHUNK ? left : AG_CHUNK);
      src = peers.base[p] + off + (long long)p * slice_bytes + o;
      dst = peers.base[rank] + off + (long long)p * slice_bytes + o;
    };
    auto issue_load = [&](long long k) {
      const uint8_t* src; uint8_t* dst; uint32_t bytes;
      locate(k, src, dst, bytes);
      const int st = (int)(k % AG_STAGES);
      mbar_arrive_expect_tx(&bars[st], bytes);
      bulk_g2s(ring + st * AG_CHUNK, src, bytes, &bars[st]);
    };
    for (long long k = 0; k < AG_AHEAD && k < n; ++k) issue_load(k);
    for (long long k = 0; k < n; ++k) {
      if (k + AG_AHEAD < n) {
        asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(AG_STAGES - AG_AHEAD - 1) : "memory");
        issue_load(k + AG_AHEAD);
      }
      const int st = (int)(k % AG_STAGES);
      mbar_wait(&bars[st], (uint32_t)((k / AG_STAGES) & 1));
      const uint8_t* src; uint8_t* dst; uint32_t bytes;
      locate(k, src, dst, bytes);
      bulk_s2g(dst, ring + st * AG_CHUNK, bytes);
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
  __syncwarp();
  if (last_cta(counter)) {
    __threadfence_system();
    signal_all(peers, rank, world, ch_done, epoch);
    wait_all(peers, rank, world, ch_done, epoch);
  }
}

// ------------------------------------------------------------------------------ reduce-scatter + AdamW
struct AdamArgs {
  float lr, beta1, beta2, eps, weight_decay, bias_c1, bias_c2, grad_scale;
};

// Rank r owns elements [r*n, (r+1)*n) of every rank's flat bf16 gradient (at byte offset grad_off of
// the symmetric buffers).  It sums the copies of all ranks, applies AdamW to its fp32 master /
// moments and writes the new bf16 parameter shard (symmetric memory at param_off, where peers
// all-gather it from).  do_adam == 0: only the reduced fp32 gradient shard is written (grad_out).
//
// Data path: warp 0 streams tiles of 1024 elements from every peer into shared memory with
// cp.async.bulk (TMA engine, NVLink), warps 1-4 reduce them out of shared memory and run the
// optimizer math; a full/empty mbarrier pair per stage overlaps the two.
static constexpr int RS_TILE = 1024;           // elements per tile (128 consumer threads x 8)
static constexpr int RS_TILE_BYTES = RS_TILE * 2;

// Element-unit segment table for the reduce-scatter: piece i of my shard covers shard elements [start, start + len) and is
// the sum over ranks of gradient elements [full + slice * len, full + (slice + 1) * len) of every rank's flat gradient
// (slice = the index of the pieces I own: my rank, or my rank inside the shard group under hybrid sharding, where the
// `world` ranks being summed are shard x replicate).  Tiles are enumerated per segment, so len only has to be a multiple of 8.
struct RsSegs {
  long long start[MAX_SEGS], full[MAX_SEGS], len[MAX_SEGS];
  int n, slice;
};
__device__ __forceinline__ long long rs_total_tiles(const RsSegs& sg) {
  long long t = 0;
  for (int i = 0; i < sg.n; ++i) t += (sg.len[i] + RS_TILE - 1) / RS_TILE;
  return t;
}
// tile t -> (segment, first element inside the segment, elements in the tile)
__device__ __forceinline__ void rs_locate(const RsSegs& sg, long long t, int& si, long long& in_seg, int& cnt) {
  si = 0;
  for (; si + 1 < sg.n; ++si) {
    const long long c = (sg.len[si] + RS_TILE - 1) / RS_TILE;
    if (t < c) break;
    t -= c;
  }
  in_seg = t * RS_TILE;
  const long long left = sg.len[si] - in_seg;
  cnt = (int)(left < RS_TILE ? left : RS_TILE);
}

__global__ void __launch_bounds__(160)
reduce_scatter_adamw_kernel(Peers peers, int rank, int world, long long grad_off, long long n, RsSegs segs,
                            float* __restrict__ master, float* __restrict__ m, float* __restrict__ v,
                            long long param_off, float* __restrict__ grad_out, AdamArgs a, int do_adam,
                            const float* __restrict__ clip_coef, int ch_arrive, int ch_done, uint32_t epoch,
                            uint32_t* __restrict__ counter, int stages) {
  extern __shared__ uint8_t rs_smem_raw[];
  uint8_t* ring = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(rs_smem_raw) + 127) & ~(uintptr_t)127);
  const int stage_bytes = world * RS_TILE_BYTES;
  uint64_t* full = reinterpret_cast<uint64_t*>(ring + stages * stage_bytes);
  uint64_t* empty = full + 4;
  if (threadIdx.x == 0) {
    for (int i = 0; i < stages; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 4);
    }
    fence_barrier_init();
  }
  // (1) everybody's gradients are complete
  if (blockIdx.x == 0) signal_all(peers, rank, world, ch_arrive, epoch);
  wait_all(peers, rank, world, ch_arrive, epoch);  // ends with __syncthreads(): barrier inits are visible

  const long long ntiles = rs_total_tiles(segs);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0) {
    if (lane == 0) {
      int st = 0;
      uint32_t ph = 0;
      for (long long t = blockIdx.x; t < ntiles; t += gridDim.x) {
        mbar_wait(&empty[st], ph ^ 1);
        int si, cnt;
        long long in_seg;
        rs_locate(segs, t, si, in_seg, cnt);
        const uint32_t bytes = (uint32_t)(cnt * 2);
        const long long src_elem = segs.full[si] + (long long)segs.slice * segs.len[si] + in_seg;
        mbar_arrive_expect_tx(&full[st], bytes * world);
        for (int s = 0; s < world; ++s) {
          const int p = (rank + s) % world;
          bulk_g2s(ring + st * stage_bytes + p * RS_TILE_BYTES, peers.base[p] + grad_off + (src_elem << 1), bytes, &full[st]);
        }
        if (++st == stages) { st = 0; ph ^= 1; }
      }
    }
  } else {
    const int ct = threadIdx.x - 32;  // 0..127
    const float gs = a.grad_scale * (clip_coef ? *clip_coef : 1.f);
    __nv_bfloat16* shard = reinterpret_cast<__nv_bfloat16*>(peers.base[rank] + param_off);
    int st = 0;
    uint32_t ph = 0;
    for (long long t = blockIdx.x; t < ntiles; t += gridDim.x) {
      mbar_wait(&full[st], ph);
      int si, cnt;
      long long in_seg;
      rs_locate(segs, t, si, in_seg, cnt);
      const bool valid = ct * 8 < cnt;
      const long long e0 = valid ? segs.start[si] + in_seg + ct * 8 : n;
      float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      if (e0 < n) {
        for (int p = 0; p < world; ++p) {
          const uint4 vv = *reinterpret_cast<const uint4*>(ring + st * stage_bytes + p * RS_TILE_BYTES + ct * 16);
          accum_bf16x8(acc, vv);
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty[st]);  // the stage may be refilled while we do the optimizer math
      if (++st == stages) { st = 0; ph ^= 1; }
      if (e0 >= n) continue;
      if (do_adam == 1) {
        float pw[8];
        const float4 m0 = *reinterpret_cast<const float4*>(m + e0), m1 = *reinterpret_cast<const float4*>(m + e0 + 4);
        const float4 v0 = *reinterpret_cast<const float4*>(v + e0), v1 = *reinterpret_cast<const float4*>(v + e0 + 4);
        const float4 p0 = *reinterpret_cast<const float4*>(master + e0), p1 = *reinterpret_cast<const float4*>(master + e0 + 4);
        float mm[8] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w};
        float vv[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
        float pp[8] = {p0.x, p0.y, p0.z, p0.w, p1.x, p1.y, p1.z, p1.w};
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const float g = acc[k] * gs;
          mm[k] = mm[k] * a.beta1 + (1.f - a.beta1) * g;
          vv[k] = vv[k] * a.beta2 + (1.f - a.beta2) * g * g;
          float p = pp[k] * (1.f - a.lr * a.weight_decay);
          p -= (a.lr / a.bias_c1) * (mm[k] / (sqrtf(vv[k] / a.bias_c2) + a.eps));
          pw[k] = p;
        }
        *reinterpret_cast<float4*>(m + e0) = make_float4(mm[0], mm[1], mm[2], mm[3]);
        *reinterpret_cast<float4*>(m + e0 + 4) = make_float4(mm[4], mm[5], mm[6], mm[7]);
        *reinterpret_cast<float4*>(v + e0) = make_float4(vv[0], vv[1], vv[2], vv[3]);
        *reinterpret_cast<float4*>(v + e0 + 4) = make_float4(vv[4], vv[5], vv[6], vv[7]);
        *reinterpret_cast<float4*>(master + e0) = make_float4(pw[0], pw[1], pw[2], pw[3]);
        *reinterpret_cast<float4*>(master + e0 + 4) = make_float4(pw[4], pw[5], pw[6], pw[7]);
        uint4 o;
        o.x = f2_to_bf2(pw[0], pw[1]); o.y = f2_to_bf2(pw[2], pw[3]);
        o.z = f2_to_bf2(pw[4], pw[5]); o.w = f2_to_bf2(pw[6], pw[7]);
        *reinterpret_cast<uint4*>(shard + e0) = o;
      } else if (do_adam == 2) {
        // all-reduce phase 1: the scaled sum goes back, as bf16, into my slice of my own buffer (param_off points there)
        uint4 o;
        o.x = f2_to_bf2(acc[0] * gs, acc[1] * gs); o.y = f2_to_bf2(acc[2] * gs, acc[3] * gs);
        o.z = f2_to_bf2(acc[4] * gs, acc[5] * gs); o.w = f2_to_bf2(acc[6] * gs, acc[7] * gs);
        *reinterpret_cast<uint4*>(shard + e0) = o;
      } else {
        *reinterpret_cast<float4*>(grad_out + e0) = make_float4(acc[0] * gs, acc[1] * gs, acc[2] * gs, acc[3] * gs);
        *reinterpret_cast<float4*>(grad_out + e0 + 4) = make_float4(acc[4] * gs, acc[5] * gs, acc[6] * gs, acc[7] * gs);
      }
    }
  }
  // (2) my shard is updated and I no longer read anybody's gradients; when every peer says the same,
  //     my own gradient buffer is free again, which is what kernel completion promises the stream.
  if (last_cta(counter)) {
    __threadfence_system();
    signal_all(peers, rank, world, ch_done, epoch);
    wait_all(peers, rank, world, ch_done, epoch);
  }
}

// ------------------------------------------------------------------------------ NVLS (in-switch) variants
// When the symmetric buffers are bound to an NVLink multicast object (parallel/symm.py takes the mapping from
// torch.distributed._symmetric_memory), `mc` is the multicast alias of the same buffer: a multimem.ld_reduce through it
// returns the SUM of the addressed bytes of every rank - the NVSwitch adds them - and a multimem.st stores to every rank.
// Per reduced byte the GPU then receives 1 stream instead of `world`, and nothing is staged in shared memory.
__device__ __forceinline__ uint4 multimem_ld_reduce_bf16x8(const void* mc) {
  uint4 r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(mc)
               : "memory");
  return r;
}
__device__ __forceinline__ void multimem_st_bf16x8(void* mc, const uint4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.bf16x2 [%0], {%1, %2, %3, %4};" ::"l"(mc), "r"(v.x), "r"(v.y), "r"(v.z),
               "r"(v.w)
               : "memory");
}

// Reduce-scatter + AdamW with the reduction done by the switch.  Same contract as reduce_scatter_adamw_kernel; every thread
// owns 8 consecutive shard elements per iteration: one multimem.ld_reduce (16 B) brings their sum over all ranks.
static constexpr int NVLS_UNROLL = 4;
__global__ void __launch_bounds__(512)
reduce_scatter_adamw_nvls_kernel(Peers peers, const uint8_t* __restrict__ mc, int rank, int world, long long grad_off, long long n,
                                 RsSegs segs, float* __restrict__ master, float* __restrict__ m, float* __restrict__ v,
                                 long long param_off, float* __restrict__ grad_out, AdamArgs a, int do_adam,
                                 const float* __restrict__ clip_coef, int ch_arrive, int ch_done, uint32_t epoch,
                                 uint32_t* __restrict__ counter) {
  if (blockIdx.x == 0) signal_all(peers, rank, world, ch_arrive, epoch);
  wait_all(peers, rank, world, ch_arrive, epoch);
  const float gs = a.grad_scale * (clip_coef ? *clip_coef : 1.f);
  __nv_bfloat16* shard = reinterpret_cast<__nv_bfloat16*>(peers.base[rank] + param_off);
  const long long nvec = n >> 3;  // every segment length is a multiple of 8
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i0 = (long long)blockIdx.x * blockDim.x + threadIdx.x; i0 < nvec; i0 += stride * NVLS_UNROLL) {
    uint4 sums[NVLS_UNROLL];
    long long e[NVLS_UNROLL];
#pragma unroll
    for (int u = 0; u < NVLS_UNROLL; ++u) {
      const long long i = i0 + u * stride;
      e[u] = -1;
      if (i < nvec) {
        const long long e0 = i << 3;
        int si = 0;
        while (si + 1 < segs.n && e0 >= segs.start[si] + segs.len[si]) ++si;
        const long long src = segs.full[si] + (long long)segs.slice * segs.len[si] + (e0 - segs.start[si]);
        sums[u] = multimem_ld_reduce_bf16x8(mc + grad_off + (src << 1));
        e[u] = e0;
      }
    }
#pragma unroll
    for (int u = 0; u < NVLS_UNROLL; ++u) {
      if (e[u] < 0) continue;
      const long long e0 = e[u];
      float g[8];
      float2 t;
      t = bf2_to_f2(sums[u].x); g[0] = t.x * gs; g[1] = t.y * gs;
      t = bf2_to_f2(sums[u].y); g[2] = t.x * gs; g[3] = t.y * gs;
      t = bf2_to_f2(sums[u].z); g[4] = t.x * gs; g[5] = t.y * gs;
      t = bf2_to_f2(sums[u].w); g[6] = t.x * gs; g[7] = t.y * gs;
      if (do_adam == 1) {
        const float4 m0 = *reinterpret_cast<const float4*>(m + e0), m1 = *reinterpret_cast<const float4*>(m + e0 + 4);
        const float4 v0 = *reinterpret_cast<const float4*>(v + e0), v1 = *reinterpret_cast<const float4*>(v + e0 + 4);
        const float4 p0 = *reinterpret_cast<const float4*>(master + e0), p1 = *reinterpret_cast<const float4*>(master + e0 + 4);
        float mm[8] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w};
        float vv[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
        float pp[8] = {p0.x, p0.y, p0.z, p0.w, p1.x, p1.y, p1.z, p1.w};
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          mm[k] = mm[k] * a.beta1 + (1.f - a.beta1) * g[k];
          vv[k] = vv[k] * a.beta2 + (1.f - a.beta2) * g[k] * g[k];
          float p = pp[k] * (1.f - a.lr * a.weight_decay);
          p -= (a.lr / a.bias_c1) * (mm[k] / (sqrtf(vv[k] / a.bias_c2) + a.eps));
          pp[k] = p;
        }
        *reinterpret_cast<float4*>(m + e0) = make_float4(mm[0], mm[1], mm[2], mm[3]);
        *reinterpret_cast<float4*>(m + e0 + 4) = make_float4(mm[4], mm[5], mm[6], mm[7]);
        *reinterpret_cast<float4*>(v + e0) = make_float4(vv[0], vv[1], vv[2], vv[3]);
        *reinterpret_cast<float4*>(v + e0 + 4) = make_float4(vv[4], vv[5], vv[6], vv[7]);
        *reinterpret_cast<float4*>(master + e0) = make_float4(pp[0], pp[1], pp[2], pp[3]);
        *reinterpret_cast<float4*>(master + e0 + 4) = make_float4(pp[4], pp[5], pp[6], pp[7]);
        uint4 o;
        o.x = f2_to_bf2(pp[0], pp[1]); o.y = f2_to_bf2(pp[2], pp[3]);
        o.z = f2_to_bf2(pp[4], pp[5]); o.w = f2_to_bf2(pp[6], pp[7]);
        *reinterpret_cast<uint4*>(shard + e0) = o;
      } else {
        *reinterpret_cast<float4*>(grad_out + e0) = make_float4(g[0], g[1], g[2], g[3]);
        *reinterpret_cast<float4*>(grad_out + e0 + 4) = make_float4(g[4], g[5], g[6], g[7]);
      }
    }
  }
  if (last_cta(counter)) {
    __threadfence_system();
    signal_all(peers, rank, world, ch_done, epoch);
    wait_all(peers, rank, world, ch_done, epoch);
  }
}

// One-pass in-place bf16 all-reduce: rank r sums slice r of every rank inside the switch (multimem.ld_reduce), scales it and
// writes it back to every rank with one multicast store (multimem.st).  n % (8 * world) == 0.
template <int UNROLL>
__global__ void __launch_bounds__(512)
allreduce_nvls_kernel(Peers peers, uint8_t* __restrict__ mc, int rank, int world, long long off, long long n, float scale,
                      int ch_arrive, int ch_done, uint32_t epoch, uint32_t* __restrict__ counter) {
  if (blockIdx.x == 0) signal_all(peers, rank, world, ch_arrive, epoch);
  wait_all(peers, rank, world, ch_arrive, epoch);
  const long long nvec = (n >> 3) / world;
  const long long stride = (long long)gridDim.x * blockDim.x;
  uint8_t* base = mc + off + (((long long)rank * nvec) << 4);
  for (long long i0 = (long long)blockIdx.x * blockDim.x + threadIdx.x; i0 < nvec; i0 += stride * UNROLL) {
    uint4 sums[UNROLL];
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const long long i = i0 + u * stride;
      if (i < nvec) sums[u] = multimem_ld_reduce_bf16x8(base + (i << 4));
    }
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const long long i = i0 + u * stride;
      if (i >= nvec) continue;
      uint4 o = sums[u];
      if (scale != 1.f) {
        float2 t;
        t = bf2_to_f2(o.x); o.x = f2_to_bf2(t.x * scale, t.y * scale);
        t = bf2_to_f2(o.y); o.y = f2_to_bf2(t.x * scale, t.y * scale);
        t = bf2_to_f2(o.z); o.z = f2_to_bf2(t.x * scale, t.y * scale);
        t = bf2_to_f2(o.w); o.w = f2_to_bf2(t.x * scale, t.y * scale);
      }
      multimem_st_bf16x8(base + (i << 4), o);
    }
  }
  if (last_cta(counter)) {
    __threadfence_system();
    signal_all(peers, rank, world, ch_done, epoch);
    wait_all(peers, rank, world, ch_done, epoch);  // every rank's stores to my buffer have been issued and fenced
  }
}

// ------------------------------------------------------------------------------ all-reduce (two-shot, in place)
// buf (at byte offset `off` of every rank's symmetric buffer) holds n elements.  Rank r reduces
// slice r of all ranks, scales it, and stores the result back into slice r of all ranks.
template <typename T>
__global__ void __launch_bounds__(512)
allreduce_kernel(Peers peers, int rank, int world, long long off, long long n, float scale, int ch_arrive,
                 int ch_done, uint32_t epoch, uint32_t* __restrict__ counter) {
  if (blockIdx.x == 0) signal_all(peers, rank, world, ch_arrive, epoch);
  wait_all(peers, rank, world, ch_arrive, epoch);
  constexpr int EPV = 16 / sizeof(T);
  const long long nvec_total = n / EPV;  // n is padded to world * EPV by the host
  const long long nvec = nvec_total / world;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += stride) {
    const long long byte = off + (((long long)rank * nvec + i) << 4);
    uint4 vals[MAX_RANKS];
#pragma unroll
    for (int s = 0; s < MAX_RANKS; ++s)
      if (s < world) vals[s] = ld_peer_v4(peers.base[(rank + s) % world] + byte);
    uint4 o;
    if (sizeof(T) == 2) {
      float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int s = 0; s < MAX_RANKS; ++s)
        if (s < world) accum_bf16x8(acc, vals[s]);
      o.x = f2_to_bf2(acc[0] * scale, acc[1] * scale); o.y = f2_to_bf2(acc[2] * scale, acc[3] * scale);
      o.z = f2_to_bf2(acc[4] * scale, acc[5] * scale); o.w = f2_to_bf2(acc[6] * scale, acc[7] * scale);
    } else {
      float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int s = 0; s < MAX_RANKS; ++s)
        if (s < world) {
          acc[0] += __uint_as_float(vals[s].x); acc[1] += __uint_as_float(vals[s].y);
          acc[2] += __uint_as_float(vals[s].z); acc[3] += __uint_as_float(vals[s].w);
        }
      o.x = __float_as_uint(acc[0] * scale); o.y = __float_as_uint(acc[1] * scale);
      o.z = __float_as_uint(acc[2] * scale); o.w = __float_as_uint(acc[3] * scale);
    }
#pragma unroll
    for (int s = 0; s < MAX_RANKS; ++s)
      if (s < world) {
        uint8_t* dst = peers.base[(rank + s) % world] + byte;
        asm volatile("st.global.relaxed.sys.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(dst), "r"(o.x), "r"(o.y),
                     "r"(o.z), "r"(o.w)
                     : "memory");
      }
  }
  if (last_cta(counter)) {
    __threadfence_system();
    signal_all(peers, rank, world, ch_done, epoch);
    wait_all(peers, rank, world, ch_done, epoch);  // every slice of my buffer now holds the result
  }
}

// ------------------------------------------------------------------------------ all-to-all (pull)
// Every rank holds `world` send blocks of block_bytes each at send_off (block d is destined to rank
// d).  Rank r pulls block r of every peer p into out[p * block_bytes].
__global__ void __launch_bounds__(512)
alltoall_pull_kernel(Peers peers, int rank, int world, long long send_off, long long block_bytes,
                     uint8_t* __restrict__ out, int ch_arrive, int ch_done, uint32_t epoch,
                     uint32_t* __restrict__ counter) {
  if (blockIdx.x == 0) signal_all(peers, rank, world, ch_arrive, epoch);
  wait_all(peers, rank, world, ch_arrive, epoch);
  const long long nvec = block_bytes >> 4;
  const long long total = nvec * world;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int slot = (int)(i / nvec);
    const long long j = i - (long long)slot * nvec;
    const int p = (rank + 1 + slot) % world;
    const uint4 v = ld_peer_v4(peers.base[p] + send_off + (long long)rank * block_bytes + (j << 4));
    *reinterpret_cast<uint4*>(out + (long long)p * block_bytes + (j << 4)) = v;
  }
  if (last_cta(counter)) {
    __threadfence_system();
    signal_all(peers, rank, world, ch_done, epoch);
    wait_all(peers, rank, world, ch_done, epoch);  // my send blocks may be overwritten after this kernel
  }
}

// ------------------------------------------------------------------------------ Ulysses sequence <-> head exchange
// The all-to-all of Ulysses sequence parallelism, with both layout changes folded into the copy (C6): the reference
// path (DeepSpeed UlyssesSPAttentionHF) permutes to a contiguous send buffer, calls all_to_all_single and permutes
// again.  Here every rank PULLS, row by row, exactly the column segments it needs from each peer's matrix and drops
// them at their final place, with TMA bulk copies through a small smem ring (same engine-driven scheme as the
// all-gather above, so the SMs stay free and the kernel co-resides with a GEMM).
//   mode 0 (sequence-sharded -> head-sharded): peer p holds [B*S_loc, src_pitch]; I take, for every segment, the
//          `bytes`-wide column block number `rank` and write it to row (b*S + p*S_loc + r) of my [B*S, dst_pitch].
//   mode 1 (head-sharded -> sequence-sharded): peer p holds [B*S, src_pitch]; I take rows (b*S + rank*S_loc + r) and
//          write each segment to column block number p of my [B*S_loc, dst_pitch].
struct ExSeg {
  long long src_base, dst_base, bytes;
};
struct ExSegs {
  ExSeg s[3];
  int n;
};
// One thread per CTA drives the copies, so throughput = bytes in flight / round-trip time.  The ring is EX_RING bytes whatever the
// row segments look like: CHUNK is the largest segment (rounded up to 512 B .. 4 KB), the ring holds EX_RING / CHUNK stages and all but
// two of them are in flight — with the 1 KB segments of sp=8 that is 22 copies per CTA instead of 4, and several CTAs share an SM.
static constexpr int EX_RING = 24576;
static constexpr int EX_SMEM = EX_RING + 512 + 128;

template <int EX_CHUNK>
__global__ void __launch_bounds__(32)
seq_head_exchange_kernel(Peers peers, int rank, int world, long long src_off, uint8_t* __restrict__ out, int mode, int B,
                         int S_loc, long long src_pitch, long long dst_pitch, ExSegs segs, int ch_arrive, int ch_done,
                         uint32_t epoch, uint32_t* __restrict__ counter) {
  extern __shared__ uint8_t ex_smem_raw[];
  uint8_t* ring = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(ex_smem_raw) + 127) & ~(uintptr_t)127);
  constexpr int EX_STAGES = EX_RING / EX_CHUNK;
  constexpr int EX_AHEAD = EX_STAGES - 2;
  uint64_t* bars = reinterpret_cast<uint64_t*>(ring + EX_RING);
  if (blockIdx.x == 0) signal_all(peers, rank, world, ch_arrive, epoch);  // my source matrix is complete (stream order)
  wait_all(peers, rank, world, ch_arrive, epoch);
  if (threadIdx.x == 0) {
    for (int i = 0; i < EX_STAGES; ++i) mbar_init(&bars[i], 1);
    fence_barrier_init();
  }
  __syncwarp();
  if (threadIdx.x == 0) {
    int seg_chunks[3], chunks_per_row = 0;
    for (int i = 0; i < segs.n; ++i) {
      seg_chunks[i] = (int)((segs.s[i].bytes + EX_CHUNK - 1) / EX_CHUNK);
      chunks_per_row += seg_chunks[i];
    }
    const long long rows = (long long)B * S_loc;
    const long long per_peer = rows * chunks_per_row;
    const long long total = per_peer * world;
    const long long S = (long long)S_loc * world;
    const long long n = (total > blockIdx.x) ? (total - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    auto locate = [&](long long k, const uint8_t*& src, uint8_t*& dst, uint32_t& bytes) {
      const long long g = blockIdx.x + k * (long long)gridDim.x;
      const int slot = (int)(g / per_peer);
      long long rem = g - (long long)slot * per_peer;
      const long long row = rem / chunks_per_row;
      int c = (int)(rem - row * chunks_per_row);
      int si = 0;
      while (c >= seg_chunks[si]) { c -= seg_chunks[si]; ++si; }
      const ExSeg sg = segs.s[si];
      const long long o = (long long)c * EX_CHUNK;
      const long long left = sg.bytes - o;
      bytes = (uint32_t)(left < EX_CHUNK ? left : EX_CHUNK);
      const int p = (rank + 1 + slot) % world;
      const long long b = row / S_loc, r = row - b * S_loc;
      if (mode == 0) {
        src = peers.base[p] + src_off + row * src_pitch + sg.src_base + (long long)rank * sg.bytes + o;
        dst = out + (b * S + (long long)p * S_loc + r) * dst_pitch + sg.dst_base + o;
      } else {
        src = peers.base[p] + src_off + (b * S + (long long)rank * S_loc + r) * src_pitch + sg.src_base + o;
        dst = out + row * dst_pitch + sg.dst_base + (long long)p * sg.bytes + o;
      }
    };
    auto issue_load = [&](long long k) {
      const uint8_t* src; uint8_t* dst; uint32_t bytes;
      locate(k, src, dst, bytes);
      const int st = (int)(k % EX_STAGES);
      mbar_arrive_expect_tx(&bars[st], bytes);
      bulk_g2s(ring + st * EX_CHUNK, src, bytes, &bars[st]);
    };
    for (long long k = 0; k < EX_AHEAD && k < n; ++k) issue_load(k);
    for (long long k = 0; k < n; ++k) {
      if (k + EX_AHEAD < n) {
        asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(EX_STAGES - EX_AHEAD - 1) : "memory");
        issue_load(k + EX_AHEAD);
      }
      const int st = (int)(k % EX_STAGES);
      mbar_wait(&bars[st], (uint32_t)((k / EX_STAGES) & 1));
      const uint8_t* src; uint8_t* dst; uint32_t bytes;
      locate(k, src, dst, bytes);
      bulk_s2g(dst, ring + st * EX_CHUNK, bytes);
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
  __syncwarp();
  if (last_cta(counter)) {
    __threadfence_system();
    signal_all(peers, rank, world, ch_done, epoch);
    wait_all(peers, rank, world, ch_done, epoch);  // every peer finished reading my source: it may be reused
  }
}

static inline int load_peers(Peers& P, const void* const* peer_ptrs, int world) {
  if (world < 1 || world > MAX_RANKS) return B200_ERR_UNSUPPORTED;
  for (int i = 0; i < MAX_RANKS; ++i) P.base[i] = i < world ? (uint8_t*)peer_ptrs[i] : nullptr;
  return 0;
}

}  // namespace b200

using namespace b200;

B200_API int b200_comm_limits(int* max_ranks, int* num_channels, long long* signal_pad_bytes) {
  if (max_ranks) *max_ranks = MAX_RANKS;
  if (num_channels) *num_channels = NUM_CHANNELS;
  if (signal_pad_bytes) *signal_pad_bytes = (long long)NUM_CHANNELS * MAX_RANKS * 4;
  return 0;
}

B200_API int b200_comm_set_timeout_ms(long long ms) {
  unsigned long long ns = (unsigned long long)ms * 1000000ull;
  B200_CUDA(cudaMemcpyToSymbol(g_wait_timeout_ns, &ns, sizeof(ns)));
  return 0;
}

// peer_ptrs: HOST array of `world` device pointers (base of each rank's symmetric buffer)
B200_API int b200_barrier_all(const void* const* peer_ptrs, int rank, int world, int channel, unsigned epoch,
                              void* stream) {
  Peers P;
  int rc = load_peers(P, peer_ptrs, world);
  if (rc) return rc;
  barrier_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(P, rank, world, channel, epoch);
  b200_count_launch();
  B200_CHECK_LAUNCH();
  return 0;
}

B200_API int b200_allgather_pull(const void* const* peer_ptrs, int rank, int world, long long shard_off,
                                 long long shard_bytes, void* out, int wait_channel, unsigned wait_epoch,
                                 int ctas, void* stream) {
  if (shard_bytes % 16 || shard_off % 16) return B200_ERR_BAD_ARG;
  Peers P;
  int rc = load_peers(P, peer_ptrs, world);
  if (rc) return rc;
  if (ctas <= 0) ctas = 32;
  static bool attr = false;
  if (!attr) {
    B200_CUDA(cudaFuncSetAttribute(allgather_pull_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, AG_SMEM));
    attr = true;
  }
  Segs sg;
  memset(&sg, 0, sizeof(sg));
  sg.n = 1;
  sg.bytes[0] = shard_bytes;
  allgather_pull_kernel<<<ctas, 32, AG_SMEM, (cudaStream_t)stream>>>(P, rank, world, shard_off, sg,
                                                                     (uint8_t*)out, wait_channel, wait_epoch);
  b200_count_launch();
  B200_CHECK_LAUNCH();
  return 0;
}

// All-gather of a sharded unit with a segment table: segs = host array of nseg x {shard_off, full_off, bytes} (bytes,
// multiples of 16; shard_off relative to shard_off_base).  Rank q's piece of segment i lands at out + full_off + q * bytes.
B200_API int b200_allgather_pull_segs(const void* const* peer_ptrs, int rank, int world, long long shard_off_base,
                                      const long long* segs, int nseg, void* out, int wait_channel, unsigned wait_epoch,
                                      int ctas, void* stream) {
  if (nseg < 1 || nseg > MAX_SEGS || shard_off_base % 16) return B200_ERR_BAD_ARG;
  Peers P;
  int rc = load_peers(P, peer_ptrs, world);
  if (rc) return rc;
  Segs sg;
  memset(&sg, 0, sizeof(sg));
  sg.n = nseg;
  for (int i = 0; i < nseg; ++i) {
    sg.shard_off[i] = segs[3 * i]; sg.full_off[i] = segs[3 * i + 1]; sg.bytes[i] = segs[3 * i + 2];
    if ((sg.shard_off[i] | sg.full_off[i] | sg.bytes[i]) % 16 || sg.bytes[i] <= 0) return B200_ERR_BAD_ARG;
  }
  if (ctas <= 0) ctas = 32;
  static bool attr = false;
  if (!attr) {
    B200_CUDA(cudaFuncSetAttribute(allgather_pull_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, AG_SMEM));
    attr = true;
  }
  allgather_pull_kernel<<<ctas, 32, AG_SMEM, (cudaStream_t)stream>>>(P, rank, world, shard_off_base, sg, (uint8_t*)out,
                                                                     wait_channel, wait_epoch);
  b200_count_launch();
  B200_CHECK_LAUNCH();
  return 0;
}

// n = elements in MY shard (multiple of 8); every rank's flat gradient holds world * n bf16 elements
B200_API int b200_reduce_scatter_adamw(const void* const* peer_ptrs, int rank, int world, long long grad_off,
                                       long long n, void* master, void* m, void* v, long long param_off,
                                       void* grad_out, float lr, float beta1, float beta2, float eps,
                                       float weight_decay, int step, float grad_scale, int do_adam,
                                       const void* clip_coef, int ch_arrive, int ch_done, unsigned epoch,
                                       void* counter, int ctas, void* stream) {
  if (n % 8 || grad_off % 16 || param_off % 16) return B200_ERR_BAD_ARG;
  Peers P;
  int rc = load_peers(P, peer_ptrs, world);
  if (rc) return rc;
  AdamArgs a{lr, beta1, beta2, eps, weight_decay, 1.f - powf(beta1, (float)step), 1.f - powf(beta2, (float)step),
             grad_scale};
  if (ctas <= 0) ctas = 148;
  int stages = 24576 / (world * RS_TILE_BYTES);
  stages = stages < 1 ? 1 : (stages > 3 ? 3 : stages);
  const int smem = stages * world * RS_TILE_BYTES + 128 + 128;
  static bool attr = false;
  if (!attr) {
    B200_CUDA(cudaFuncSetAttribute(reduce_scatter_adamw_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 3 * 8 * RS_TILE_BYTES + 256));
    attr = true;
  }
  RsSegs sg;
  memset(&sg, 0, sizeof(sg));
  sg.n = 1;
  sg.len[0] = n;
  sg.slice = rank;
  reduce_scatter_adamw_kernel<<<ctas, 160, smem, (cudaStream_t)stream>>>(
      P, rank, world, grad_off, n, sg, (float*)master, (float*)m, (float*)v, param_off, (float*)grad_out, a, do_adam,
      (const float*)clip_coef, ch_arrive, ch_done, epoch, (uint32_t*)counter, stages);
  b200_count_launch();
  B200_CHECK_LAUNCH();
  return 0;
}

// Same kernel for a unit whose shard is made of several pieces (see RsSegs): segs = host array of nseg x {start, full, len}
// in ELEMENTS; start values ascending and contiguous, every len a multiple of 8.  n = sum of len.  slice = index of the pieces
// this rank owns (== rank, except under hybrid sharding where `world` spans shard x replicate and slice is the shard rank).
B200_API int b200_reduce_scatter_adamw_segs(const void* const* peer_ptrs, int rank, int world, int slice, long long grad_off,
                                            const long long* segs, int nseg, void* master, void* m, void* v, long long param_off,
                                            void* grad_out, float lr, float beta1, float beta2, float eps, float weight_decay,
                                            int step, float grad_scale, int do_adam, const void* clip_coef, int ch_arrive,
                                            int ch_done, unsigned epoch, void* counter, int ctas, void* stream) {
  if (nseg < 1 || nseg > MAX_SEGS || grad_off % 16 || param_off % 16) return B200_ERR_BAD_ARG;
  Peers P;
  int rc = load_peers(P, peer_ptrs, world);
  if (rc) return rc;
  RsSegs sg;
  memset(&sg, 0, sizeof(sg));
  sg.n = nseg;
  sg.slice = slice;
  long long n = 0;
  for (int i = 0; i < nseg; ++i) {
    sg.start[i] = segs[3 * i]; sg.full[i] = segs[3 * i + 1]; sg.len[i] = segs[3 * i + 2];
    if (sg.start[i] != n || sg.len[i] <= 0 || (sg.len[i] % 8) || (sg.full[i] % 8)) return B200_ERR_BAD_ARG;
    n += sg.len[i];
  }
  AdamArgs a{lr, beta1, beta2, eps, weight_decay, 1.f - powf(beta1, (float)step), 1.f - powf(beta2, (float)step),
             grad_scale};
  if (ctas <= 0) ctas = 148;
  int stages = 24576 / (world * RS_TILE_BYTES);
  stages = stages < 1 ? 1 : (stages > 3 ? 3 : stages);
  const int smem = stages * world * RS_TILE_BYTES + 128 + 128;
  static bool attr = false;
  if (!attr) {
    B200_CUDA(cudaFuncSetAttribute(reduce_scatter_adamw_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 3 * 8 * RS_TILE_BYTES + 256));
    attr = true;
  }
  reduce_scatter_adamw_kernel<<<ctas, 160, smem, (cudaStream_t)stream>>>(
      P, rank, world, grad_off, n, sg, (float*)master, (float*)m, (float*)v, param_off, (float*)grad_out, a, do_adam,
      (const float*)clip_coef, ch_arrive, ch_done, epoch, (uint32_t*)counter, stages);
  b200_count_launch();
  B200_CHECK_LAUNCH();
  return 0;
}

// in-place all-reduce of n elements (bf16 if is_bf16 else fp32) at byte offset `off`; n % (world * 16 / esize) == 0
B200_API int b200_allreduce(const void* const* peer_ptrs, int rank, int world, long long off, long long n,
                            int is_bf16, float scale, int ch_arrive, int ch_done, unsigned epoch, void* counter,
                            int ctas, void* stream) {
  const int epv = is_bf16 ? 8 : 4;
  if (off % 16 || n % ((long long)world * epv)) return B200_ERR_BAD_ARG;
  Peers P;
  int rc = load_peers(P, peer_ptrs, world);
  if (rc) return rc;
  if (ctas <= 0) ctas = 32;
  if (is_bf16)
    allreduce_kernel<__nv_bfloat16><<<ctas, 512, 0, (cudaStream_t)stream>>>(P, rank, world, off, n, scale, ch_arrive,
                                                                         ch_done, epoch, (uint32_t*)counter);
  else
    allreduce_kernel<float><<<ctas, 512, 0, (cudaStream_t)stream>>>(P, rank, world, off, n, scale, ch_arrive, ch_done,
                                                                  epoch, (uint32_t*)counter);
  b200_count_launch();
  B200_CHECK_LAUNCH();
  return 0;
}

B200_API int b200_alltoall_pull(const void* const* peer_ptrs, int rank, int world, long long send_off,
                                long long block_bytes, void* out, int ch_arrive, int ch_done, unsigned epoch,
                                void* counter, int ctas, void* stream) {
  if (block_bytes % 16 || send_off % 16) return B200_ERR_BAD_ARG;
  Peers P;
  int rc = load_peers(P, peer_ptrs, world);
  if (rc) return rc;
  if (ctas <= 0) ctas = 32;
  alltoall_pull_kernel<<<ctas, 512, 0, (cudaStream_t)stream>>>(P, rank, world, send_off, block_bytes, (uint8_t*)out,
                                                               ch_arrive, ch_done, epoch, (uint32_t*)counter);
  b200_count_launch();
  B200_CHECK_LAUNCH();
  return 0;
}

// segs: host array of 3*nseg long longs {src_base, dst_base, bytes}; all byte quantities multiples of 16
B200_API int b200_seq_head_exchange(const void* const* peer_ptrs, int rank, int world, long long src_off, void* out, int mode,
                                    int B, int S_loc, long long src_pitch, long long dst_pitch, const long long* segs, int nseg,
                                    int ch_arrive, int ch_done, unsigned epoch, void* counter, int ctas, void* stream) {
  if (nseg < 1 || nseg > 3 || (mode != 0 && mode != 1)) return B200_ERR_BAD_ARG;
  if ((src_off | src_pitch | dst_pitch | (long long)(uintptr_t)out) % 16) return B200_ERR_BAD_ARG;
  Peers P;
  int rc = load_peers(P, peer_ptrs, world);
  if (rc) return rc;
  ExSegs sg;
  sg.n = nseg;
  for (int i = 0; i < nseg; ++i) {
    sg.s[i] = ExSeg{segs[3 * i], segs[3 * i + 1], segs[3 * i + 2]};
    if ((sg.s[i].src_base | sg.s[i].dst_base | sg.s[i].bytes) % 16 || sg.s[i].bytes <= 0) return B200_ERR_BAD_ARG;
  }
  if (ctas <= 0) ctas = 3 * 148;  // 32 threads and 25 KB each: three per SM still leave a persistent GEMM CTA its shared memory
  long long widest = 0;
  for (int i = 0; i < nseg; ++i) widest = sg.s[i].bytes > widest ? sg.s[i].bytes : widest;
  static bool attr = false;
  if (!attr) {
    B200_CUDA(cudaFuncSetAttribute(seq_head_exchange_kernel<4096>, cudaFuncAttributeMaxDynamicSharedMemorySize, EX_SMEM));
    B200_CUDA(cudaFuncSetAttribute(seq_head_exchange_kernel<2048>, cudaFuncAttributeMaxDynamicSharedMemorySize, EX_SMEM));
    B200_CUDA(cudaFuncSetAttribute(seq_head_exchange_kernel<1024>, cudaFuncAttributeMaxDynamicSharedMemorySize, EX_SMEM));
    B200_CUDA(cudaFuncSetAttribute(seq_head_exchange_kernel<512>, cudaFuncAttributeMaxDynamicSharedMemorySize, EX_SMEM));
    attr = true;
  }
#define B200_EXCHANGE(CHUNK)                                                                                                    \
  seq_head_exchange_kernel<CHUNK><<<ctas, 32, EX_SMEM, (cudaStream_t)stream>>>(P, rank, world, src_off, (uint8_t*)out, mode, B, \
                                                                               S_loc, src_pitch, dst_pitch, sg, ch_arrive,      \
                                                                               ch_done, epoch, (uint32_t*)counter)
  if (widest <= 512) B200_EXCHANGE(512);
  else if (widest <= 1024) B200_EXCHANGE(1024);
  else if (widest <= 2048) B200_EXCHANGE(2048);
  else B200_EXCHANGE(4096);
#undef B200_EXCHANGE
  b200_count_launch();
  B200_CHECK_LAUNCH();
  return 0;
}

// In-place bf16 all-reduce (times `scale`) of n elements at byte offset `off`, driven by the TMA engine end to end:
//   phase 1  every rank reduces ITS slice: peers' copies -> smem (cp.async.bulk) -> fp32 sum -> bf16 back into its own buffer
//            (the reduce-scatter kernel in its "do_adam = 2" mode; it ends with a cross-rank barrier),
//   phase 2  every rank pulls the other slices from their owners (allgather_inplace_kernel, ends with a barrier).
// n % (8 * world) == 0.  Uses channels ch[0..2] and two counters (counter, counter + 1).
B200_API int b200_allreduce_tma(const void* const* peer_ptrs, int rank, int world, long long off, long long n, float scale,
                                int ch_arrive, int ch_mid, int ch_done, unsigned epoch, void* counter, int ctas, void* stream) {
  if (off % 16 || n % ((long long)world * 8)) return B200_ERR_BAD_ARG;
  Peers P;
  int rc = load_peers(P, peer_ptrs, world);
  if (rc) return rc;
  if (ctas <= 0) ctas = 148;
  const long long slice = n / world;
  AdamArgs a{0.f, 0.9f, 0.999f, 1e-8f, 0.f, 1.f, 1.f, scale};
  int stages = 24576 / (world * RS_TILE_BYTES);
  stages = stages < 1 ? 1 : (stages > 3 ? 3 : stages);
  const int smem = stages * world * RS_TILE_BYTES + 128 + 128;
  static bool attr = false;
  if (!attr) {
    B200_CUDA(cudaFuncSetAttribute(reduce_scatter_adamw_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 3 * 8 * RS_TILE_BYTES + 256));
    B200_CUDA(cudaFuncSetAttribute(allgather_inplace_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, AG_SMEM));
    attr = true;
  }
  // the slices of this all-reduce start at `off`: rank r owns elements [r * slice, (r + 1) * slice)
  RsSegs sg;
  memset(&sg, 0, sizeof(sg));
  sg.n = 1;
  sg.len[0] = slice;
  sg.slice = rank;
  reduce_scatter_adamw_kernel<<<ctas, 160, smem, (cudaStream_t)stream>>>(
      P, rank, world, off, slice, sg, nullptr, nullptr, nullptr, off + (long long)rank * slice * 2, nullptr, a, 2, nullptr, ch_arrive, ch_mid,
      epoch, (uint32_t*)counter, stages);
  b200_count_launch();
  B200_CHECK_LAUNCH();
  allgather_inplace_kernel<<<ctas, 32, AG_SMEM, (cudaStream_t)stream>>>(P, rank, world, off, slice * 2, ch_done, epoch,
                                                                       (uint32_t*)counter + 1);
  b200_count_launch();
  B200_CHECK_LAUNCH();
  return 0;
}

// NVLS versions (see the kernels): mc = multicast alias of the symmetric buffers (0 / NULL: not available -> B200_ERR_UNSUPPORTED)
B200_API int b200_reduce_scatter_adamw_nvls(const void* const* peer_ptrs, const void* mc, int rank, int world, int slice,
                                            long long grad_off, const long long* segs, int nseg, void* master, void* m, void* v,
                                            long long param_off, void* grad_out, float lr, float beta1, float beta2, float eps,
                                            float weight_decay, int step, float grad_scale, int do_adam, const void* clip_coef,
                                            int ch_arrive, int ch_done, unsigned epoch, void* counter, int ctas, void* stream) {
  if (mc == nullptr) return B200_ERR_UNSUPPORTED;
  if (nseg < 1 || nseg > MAX_SEGS || grad_off % 16 || param_off % 16 || (do_adam != 0 && do_adam != 1)) return B200_ERR_BAD_ARG;
  Peers P;
  int rc = load_peers(P, peer_ptrs, world);
  if (rc) return rc;
  RsSegs sg;
  memset(&sg, 0, sizeof(sg));
  sg.n = nseg;
  sg.slice = slice;
  long long n = 0;
  for (int i = 0; i < nseg; ++i) {
    sg.start[i] = segs[3 * i]; sg.full[i] = segs[3 * i + 1]; sg.len[i] = segs[3 * i + 2];
    if (sg.start[i] != n || sg.len[i] <= 0 || (sg.len[i] % 8) || (sg.full[i] % 8)) return B200_ERR_BAD_ARG;
    n += sg.len[i];
  }
  AdamArgs a{lr, beta1, beta2, eps, weight_decay, 1.f - powf(beta1, (float)step), 1.f - powf(beta2, (float)step),
             grad_scale};
  if (ctas <= 0) ctas = 148;
  reduce_scatter_adamw_nvls_kernel<<<ctas, 512, 0, (cudaStream_t)stream>>>(
      P, (const uint8_t*)mc, rank, world, grad_off, n, sg, (float*)master, (float*)m, (float*)v, param_off, (float*)grad_out, a,
      do_adam, (const float*)clip_coef, ch_arrive, ch_done, epoch, (uint32_t*)counter);
  b200_count_launch();
  B200_CHECK_LAUNCH();
  return 0;
}

B200_API int b200_allreduce_nvls(const void* const* peer_ptrs, void* mc, int rank, int world, long long off, long long n,
                                 float scale, int ch_arrive, int ch_done, unsigned epoch, void* counter, int ctas, void* stream) {
  if (mc == nullptr) return B200_ERR_UNSUPPORTED;
  if (off % 16 || n % ((long long)world * 8)) return B200_ERR_BAD_ARG;
  Peers P;
  int rc = load_peers(P, peer_ptrs, world);
  if (rc) return rc;
  if (ctas <= 0) ctas = 148;
  static int unroll = 0;
  if (unroll == 0) {
    const char* e = getenv("ACCELERATE_B200_NVLS_UNROLL");
    unroll = e ? atoi(e) : 4;
    if (unroll != 2 && unroll != 4 && unroll != 8) unroll = 4;
  }
  cudaStream_t st = (cudaStream_t)stream;
  if (unroll == 2)
    allreduce_nvls_kernel<2><<<ctas, 512, 0, st>>>(P, (uint8_t*)mc, rank, world, off, n, scale, ch_arrive, ch_done, epoch, (uint32_t*)counter);
  else if (unroll == 8)
    allreduce_nvls_kernel<8><<<ctas, 512, 0, st>>>(P, (uint8_t*)mc, rank, world, off, n, scale, ch_arrive, ch_done, epoch, (uint32_t*)counter);
  else
    allreduce_nvls_kernel<4><<<ctas, 512, 0, st>>>(P, (uint8_t*)mc, rank, world, off, n, scale, ch_arrive, ch_done, epoch, (uint32_t*)counter);
  b200_count_launch();
  B200_CHECK_LAUNCH();
  return 0;
}
