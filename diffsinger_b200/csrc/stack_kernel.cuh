// Whole-network kernel for LAUNCH-BOUND sessions: ONE launch = one complete backbone evaluation (WaveNet.forward,
// modules/backbones/wavenet.py:75-107: input projection, all L residual layers, skip sum, skip projection, output
// projection) for sessions of a few 128-frame tiles -- the batch-1 `.ds` segments of inference/ds_acoustic.py:234-244 and the
// two sequential variance samplers of modules/toplevel.py:281,302 -- which the per-layer kernels can only run as a chain of
// ~2L+5 dependent launches per evaluation (each ~6 us of launch -> first TMA -> epilogue -> store latency).
//
// Layout of the work: a CLUSTER of KB = C/64 CTAs owns one 128-frame tile; CTA r owns output channels [64r, 64r+64) of every
// GEMM (N-split), so each SM streams only 1/KB of the weights from L2, and the activations never leave the cluster:
//   * x~ (with a +-16-frame halo for the dilated conv) and the gated activation z live in SHARED MEMORY as K-major
//     SWIZZLE_128B operand tiles ([rows][64 channels] per k-block); after each epilogue every CTA broadcasts its 64-channel
//     slice (16 KB) to the other CTAs of the cluster with `cp.async.bulk.shared::cluster` (DSMEM) that completes on the
//     RECEIVER's mbarrier -- an all-gather without any cluster-wide barrier; every peer block has its own barrier and the
//     copies leave one after the other in the order the peers consume them, so the consumer MMA starts on a k-block as
//     soon as that block has landed (DSMEM moves only ~17-21 B/clk per SM: the exchange is bandwidth-bound);
//   * the three conv taps are three UMMA descriptors into the same x~ tile shifted by dil rows (as in layer_kernel.cuh);
//   * the skip sum is accumulated in TMEM across all L layers (64 fp32 columns per CTA) and never touches memory
//     ("skip accumulation kept on-chip");
//   * the step-invariant conditioner projection is hoisted: ds_session_set_cond computes cproj = W_cond*cond + b once per
//     sampler call (f16, [L][rows][2C]); the gate epilogue adds its row slice (prefetched before the accumulator wait);
//   * weights stream through a 5-stage TMA ring that runs ahead across layer boundaries (weights do not depend on data).
// Utterances longer than 128 frames span several clusters; neighbouring tiles exchange the 16 edge frames of x~ per layer
// through a small global (L2-resident) buffer of self-validating packets ({data, flag} per 8 bytes, the "LL" idea of NCCL):
// the sender's epilogue threads just store, the receiver's epilogue threads (idle while the centre taps run) poll the packets
// and write the halo rows -- point-to-point, no fence, no counter, no grid-wide barrier.
// All clusters of a launch must be co-resident (the host checks cudaOccupancyMaxActiveClusters).  Ragged batches
// (StackParams::len): every utterance has its own frame count -- validity masks, the right-neighbour flag and the early exit of
// tiles beyond the end follow it, so a short utterance in a padded batch computes exactly what it computes alone.
//
// Warp roles (384 threads): warps 0 / 3 / 2 = weight TMA producers, warp1 = MMA issuer (all four run converged with one elected
// issuing lane), warp2 also allocates TMEM, warps 4..11 = epilogue + halo receive (warp & 3 = TMEM lane quadrant, (warp - 4) >> 2 = which 32 of the CTA's 64 channels).
#pragma once
#include <cuda.h>
#include "gemm_kernel.cuh"
#include "layer_kernel.cuh"

namespace dsb {

constexpr int kStackMaxKB = 4;              // cluster size KB = C / 64: 4 (C = 256: pitch predictor, acoustic 20 x 256) or 3 (C = 192:
                                            // multi-variance predictor)
constexpr int kStackHalo = 16;              // halo frames each side = largest supported dilation
constexpr int kStackXRows = 128 + 2 * kStackHalo;
constexpr uint32_t kStackXTile = kStackXRows * 128u;   // 20480 B per k-block (1024-aligned)
constexpr uint32_t kStackZTile = 128u * 128u;          // 16384 B per k-block
constexpr uint32_t kStackStage = 16384u;               // weight ring stage: [128 rows][64 K] bf16
__host__ __device__ constexpr int stack_stages(int kb) { return kb == 4 ? 5 : 7; }  // weight-ring depth that fits beside x~ and z
constexpr int kStackProducers = 3;            // weight-issuing threads (warp 0, warp 3, warp 2)
__host__ __device__ constexpr uint32_t stack_smem_bytes(int kb) {
  return 1024u + kb * (kStackXTile + kStackZTile) + stack_stages(kb) * kStackStage + 1024u;
}
constexpr uint32_t kStackEdgeBytes = kStackHalo * 256u;  // one k-block of one edge: 16 rows x 8 chunks x 32 B (flagged packets, below)

struct StackParams {
  CUtensorMap tmW1;     // gate weights, 128-row passes (plan o_w1s): dims (3C + H, L * 2C), box (64, 128)
  CUtensorMap tmW2;     // residual half of output_projection (o_w2): dims (C, L * C), box (64, 64)
  CUtensorMap tmWsk;    // deferred-skip matrix (o_wsk): dims (L * C, C), box (64, 64)
  CUtensorMap tmWsp;    // skip_projection (o_wsp): dims (C, C), box (64, 64)
  CUtensorMap tmWhead;  // input_projection (o_whead): dims (FMp, C), box (64, 64)
  CUtensorMap tmWout;   // output_projection (o_wout): dims (C, FMr), box (64, nout)
  const float* x;       // [B][T][FM] fp32
  float* out;           // [B][T][FM] fp32
  const uint16_t* cproj;  // f16 [L][B*T][2C], 128-column passes [64 gate (x0.5, bias included) | 64 filter (bias included)]
  const float* bhead;   // [C]
  const float* dtab;    // step-embedding projection of layer 0 for this evaluation: [vec_bstride * b + c]
  const float* rtab;    // folded residual vectors (b2_l - d_l)/sqrt(2) + d_{l+1}: [vec_bstride * b + l * C + c]
  const float* bsk;     // [C] summed skip-half biases
  const float* bsp;     // [C]
  const float* bout;    // [FMr]
  int vec_bstride;
  uint8_t* edges;       // [2 parity][tiles][2 sides][KB][16 rows][8 chunks][32 B]: edge frames of x~ as FLAGGED packets: every 8
                        // bytes = {2 bf16 channels, flag}; the receiver polls the data itself (one L2 round trip, no fence, no
                        // counter).  Zeroed by the host at the start of every program run; flags are never 0.
  unsigned int flag_base;  // L * (index of this evaluation in the program run): layer seq's packets carry flag_base + seq + 1
  int T, B, tiles_t, n_tiles;
  // sampler update fused into the output epilogue (ds_session_run peephole: EVAL followed by LINCOMB): upd_dst = sum_k upd_coef[k] *
  // upd_src[k], evaluated like lincomb_kernel (fma chain from 0 in term order); a source equal to `out` is this evaluation's result
  int upd_n;
  float upd_coef[8];
  const float* upd_src[8];
  float* upd_dst;
  const int* len;       // optional [B]: frames of utterance b (ds_session_set_lengths); nullptr = every utterance has T frames
  int L, cycle, C, FM, FMp, nout;
  float inv_sqrt_L;
  int dbg;  // debug (DSB200_DBG, timing only -- results are wrong): bit 1 = no weight loads, bit 2 = no MMAs, bit 3 = no peer copies
  unsigned long long* trace;  // debug (ds_debug_set_trace): CTA 0 records %globaltimer at 10 points per layer, [L][10] words
  int trace_cap;
};

// ---------------------------------------------------------------- DSMEM / global bulk copies
__device__ __forceinline__ uint32_t mapa_shared(uint32_t addr, uint32_t cta) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(cta));
  return r;
}
// local shared memory -> the same offset in CTA `cta`; completes (complete_tx) on that CTA's barrier
__device__ __forceinline__ void bulk_copy_to_peer(uint32_t dst_local_addr, uint32_t src_addr, uint32_t bytes, uint32_t bar_local_addr,
                                                  uint32_t cta) {
  asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(mapa_shared(dst_local_addr, cta)), "r"(src_addr), "r"(bytes), "r"(mapa_shared(bar_local_addr, cta))
               : "memory");
}
__device__ __forceinline__ void bulk_copy_g2s(uint32_t dst_addr, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst_addr), "l"(reinterpret_cast<uint64_t>(src)), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }
__device__ __forceinline__ uint4 ld_volatile_v4(const void* p) {
  uint4 v;
  asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_volatile_v4(void* p, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync 1, 256;" ::: "memory"); }  // the 8 epilogue warps
__device__ __forceinline__ uint4 ld_shared_v4(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr) : "memory");
  return v;
}

// position of 16-byte chunk `j` of row `row` inside a [rows][128 B] SWIZZLE_128B tile at a 1024-aligned base
__device__ __forceinline__ uint32_t sw128_off(int row, int j) { return static_cast<uint32_t>(row) * 128u + ((j ^ (row & 7)) << 4); }

// Lean wait for the single-thread roles: every instruction on those threads is serial latency (~6 cycles each, no other warp hides
// it), so no spin counter; the bounded mbar_wait (traps instead of hanging) is used in the kDbg instantiation.
__device__ __forceinline__ void mbar_wait_lean(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred P;\n\t"
      "LAB_WAIT:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 P, [%0], %1;\n\t"
      "@P bra LAB_DONE;\n\t"
      "bra LAB_WAIT;\n\t"
      "LAB_DONE:\n\t}\n"
      ::"r"(smem_u32(bar)), "r"(parity)
      : "memory");
}

// kDbg = true: role switches (StackParams::dbg), %globaltimer trace, bounded waits.  The production instantiation carries none of
// it: the MMA issuer is ONE thread whose instruction count per weight stage is the kernel's critical path (tools/stack_trace.py:
// ~285 cycles per mbarrier operation, ~55 per tcgen05.mma issue).
template <int KB, bool kDbg>
__global__ void __cluster_dims__(KB, 1, 1) __launch_bounds__(kGemmThreads, 1)
ds_stack_kernel(const __grid_constant__ StackParams p) {
  constexpr int kStackStages = stack_stages(KB);
  const int dbg = kDbg ? p.dbg : 0;
  auto bwait = [&](uint64_t* bar, uint32_t parity) {
    if constexpr (kDbg) mbar_wait(bar, parity);
    else mbar_wait_lean(bar, parity);
  };
  const bool tracing = kDbg && p.trace != nullptr && blockIdx.x == 0;
  auto mark = [&](int l, int slot) {
    if constexpr (kDbg) {
      if (tracing && (threadIdx.x & 31) == 0 && (l * 10 + slot) < p.trace_cap) p.trace[l * 10 + slot] = globaltimer_ns();
    }
  };
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);
  uint8_t* xs = smem;                                   // KB x [160 rows][128 B]
  uint8_t* zs = xs + KB * kStackXTile;                  // KB x [128 rows][128 B]
  uint8_t* ring = zs + KB * kStackZTile;                // kStackStages x 16 KB
  uint64_t* bars = reinterpret_cast<uint64_t*>(ring + kStackStages * kStackStage);
  uint64_t* w_full = bars;
  uint64_t* w_empty = w_full + kStackStages;
  uint64_t* x_loc = w_empty + kStackStages;   // this CTA's own k-block of x~ is written (epilogue arrival)
  uint64_t* x_rem = x_loc + 1;                // [KB-1]: peer block (r + 1 + i) % KB has landed: 1 arrival (arming) + 16 KB
  uint64_t* z_loc = x_rem + (KB - 1);
  uint64_t* z_rem = z_loc + 1;                // [KB-1]
  uint64_t* acc1_full = z_rem + (KB - 1);
  uint64_t* acc2_full = acc1_full + 1;
  uint64_t* aux_full = acc2_full + 1;         // head input cast done
  uint64_t* halo_full = aux_full + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(halo_full + 1);
  constexpr uint32_t kRemBytes = kStackZTile;  // one peer block per barrier: the consumer starts on a block as soon as IT has landed

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);  // the shuffle tells the compiler the value is warp-uniform
  const int lane = threadIdx.x & 31;
  const int r = static_cast<int>(cluster_ctarank());
  const int tile = blockIdx.x / KB;
  const int b = tile / p.tiles_t;
  const int tt = tile % p.tiles_t;
  const int t0 = tt * 128;
  const int Tb = p.len != nullptr ? __ldg(p.len + b) : p.T;  // frames of this utterance (ragged batches)
  if (t0 >= Tb) return;  // tile beyond the utterance's end: the whole cluster leaves before any barrier exists; nobody waits for it
  const bool has_left = tt > 0, has_right = t0 + 128 < Tb;
  const bool has_nb = has_left || has_right;
  const int L = p.L, C = p.C;
  const int KBh = p.FMp / 64;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.tmW1);
    tma_prefetch_desc(&p.tmW2);
    tma_prefetch_desc(&p.tmWsk);
    tma_prefetch_desc(&p.tmWsp);
    tma_prefetch_desc(&p.tmWhead);
    tma_prefetch_desc(&p.tmWout);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < kStackStages; ++i) {
      mbar_init(&w_full[i], 1);
      mbar_init(&w_empty[i], 1);
    }
    mbar_init(x_loc, 1);
    mbar_init(z_loc, 1);
    for (int i = 0; i < KB - 1; ++i) {
      mbar_init(&x_rem[i], 1);
      mbar_init(&z_rem[i], 1);
    }
    mbar_init(acc1_full, 1);
    mbar_init(acc2_full, 1);
    mbar_init(aux_full, 1);
    mbar_init(halo_full, 1);
    fence_barrier_init();
    // phase 0 of the REMOTE blocks: this CTA's arrival + the 16 KB its owner will send (later phases are armed by the epilogue
    // lanes that send this CTA's own block of the same phase: publish())
    for (int i = 0; i < KB - 1; ++i) {
      if (dbg & 8) { mbar_arrive(&x_rem[i]); mbar_arrive(&z_rem[i]); }
      else { mbar_arrive_expect_tx(&x_rem[i], kRemBytes); mbar_arrive_expect_tx(&z_rem[i], kRemBytes); }
    }
  }
  // x~ starts as zeros: halo rows outside the utterance ARE the conv's zero padding and are never written
  for (uint32_t o = threadIdx.x * 16u; o < KB * kStackXTile; o += kGemmThreads * 16u) st_shared_v4(smem_u32(xs) + o, 0, 0, 0, 0);
  fence_proxy_async_smem();
  if (warp == 2) tmem_alloc<256>(tmem_slot);
  tc_fence_before();
  cluster_sync_all();  // barriers initialised and x~ zeroed in every CTA before any peer sends bytes
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tm_acc1 = tmem_base, tm_acc2 = tmem_base + 128, tm_skip = tmem_base + 192;
  // The next evaluation's launch may be scheduled NOW (programmatic dependent launch): a session of a few tiles leaves most SMs
  // idle, so its clusters set up (barriers, TMEM, the first weight stages) beside this launch and only wait -- griddepcontrol.wait in
  // the epilogue, before the first read of x -- for this one to finish.  Its start-up latency disappears from the step.
  grid_dep_launch();

  // ------------------------------------------------------------ weight producers (never wait for data: weights are constant)
  // A thread pays ~285 cycles for every mbarrier operation and bulk-copy issue (measured, tools/tma_bw_bench.cu: 16 KB stages from
  // one issuing thread = 55 GB/s, from three threads in different warps = 165 GB/s), so the stage sequence is dealt round robin to
  // kStackProducers threads; each walks the whole schedule and issues the stages whose sequence number is its own.
  // Stage order of one layer (the MMA issuer consumes in the same order): the centre tap of k-blocks r, r+1, .. (no halo needed,
  // blocks in arrival order), then taps 0 and 2 of the same k-blocks, then the 1x1 blocks.
  auto produce = [&](int me, bool lead) {  // `lead`: the one lane that issues (the rest of a converged warp only follows the waits)
    int q = 0;  // stage sequence number over the whole kernel; ring slot q % kStackStages, ring pass q / kStackStages
    // one stage: wait for the slot, issue the copies, THEN arm the barrier (complete_tx may precede expect_tx: the phase cannot
    // finish before this thread's arrival) -- the copy leaves one mbarrier operation earlier in the ring's refill round trip
    auto stage = [&](uint32_t bytes, auto&& issue) {
      const int st = q % kStackStages;
      if ((q % kStackProducers) == me) {
        bwait(&w_empty[st], ((q / kStackStages) & 1) ^ 1);
        if (lead) {
          if (dbg & 2) mbar_arrive(&w_full[st]);  // debug: barriers only, no bytes
          else {
            issue(ring + st * kStackStage, &w_full[st]);
            mbar_arrive_expect_tx(&w_full[st], bytes);
          }
        }
      }
      ++q;
    };
    for (int k = 0; k < KBh; ++k)
      stage(64 * 128, [&](uint8_t* s, uint64_t* bar) { tma_load_2d(&p.tmWhead, bar, s, k * 64, 64 * r); });
    for (int l = 0; l < L; ++l) {
      const int row1 = l * 2 * C + 128 * r;
      for (int j = 0; j < KB; ++j)
        stage(128 * 128, [&](uint8_t* s, uint64_t* bar) { tma_load_2d(&p.tmW1, bar, s, (KB + ((r + j) % KB)) * 64, row1); });
      for (int j = 0; j < KB; ++j)
        for (int tap = 0; tap < 3; tap += 2)
          stage(128 * 128, [&](uint8_t* s, uint64_t* bar) { tma_load_2d(&p.tmW1, bar, s, (tap * KB + ((r + j) % KB)) * 64, row1); });
      for (int j = 0; j < KB; ++j) {
        const int kb = (r + j) % KB;
        const bool res = l + 1 < L;
        stage((res ? 2u : 1u) * 64 * 128, [&](uint8_t* s, uint64_t* bar) {
          if (res) tma_load_2d(&p.tmW2, bar, s, kb * 64, l * C + 64 * r);
          tma_load_2d(&p.tmWsk, bar, s + 8192, l * C + kb * 64, 64 * r);
        });
      }
    }
    for (int j = 0; j < KB; ++j)
      stage(64 * 128, [&](uint8_t* s, uint64_t* bar) { tma_load_2d(&p.tmWsp, bar, s, ((r + j) % KB) * 64, 64 * r); });
    for (int j = 0; j < KB; ++j)
      stage(p.nout * 128, [&](uint8_t* s, uint64_t* bar) { tma_load_2d(&p.tmWout, bar, s, ((r + j) % KB) * 64, p.nout * r); });
  };

  if (warp == 0) {
    produce(0, elect_one());  // converged warp, one issuing lane (see the MMA issuer below)
  } else if (warp == 3) {
    produce(1, elect_one());
  } else if (warp == 1) {
    {
      // ------------------------------------------------------------ MMA issuer
      // The WHOLE warp runs this role converged (waits, stage bookkeeping and descriptor arithmetic stay in uniform registers);
      // one elected lane issues.  Running the role inside `if (lane == 0)` made ptxas wrap every tcgen05.mma / commit in a
      // vector->uniform register "waterfall" loop (ELECT / R2UR.BROADCAST / BRA.U.ANY), ~55 cycles per MMA.
      const bool lead = elect_one();
      // ONE thread on purpose.  Two issuers on alternating stages were tried (results stayed correct, but no faster -- the weight
      // ring's refill round trip, not the issuer, bounds a stage -- and the accumulation order, hence the low bits, varied from
      // run to run).
      constexpr uint32_t id128 = make_idesc_bf16_f32(128), id64 = make_idesc_bf16_f32(64);
      const uint32_t idout = make_idesc_bf16_f32(static_cast<uint32_t>(p.nout));
      int st = 0;
      uint32_t ph = 0;
      // descriptors = a constant (swizzle mode, SBO, version) + (shared address >> 4): all addresses here are < 256 KB
      const uint64_t d0 = make_kmajor_sw128_desc(0);
      const uint64_t dx = d0 + (smem_u32(xs) >> 4), dz = d0 + (smem_u32(zs) >> 4), dw = d0 + (smem_u32(ring) >> 4);
      constexpr uint32_t kXT = kStackXTile >> 4, kZT = kStackZTile >> 4, kST = kStackStage >> 4;
      // The wait on a stage's barrier is PROBED one stage early (test_wait issued before the previous stage's MMAs): the probe's
      // result latency (~285 cycles) then overlaps the MMA issue instead of adding to every stage.
      uint32_t pre = mbar_test_wait(&w_full[0], 0) ? 1u : 0u;
      auto w_wait = [&]() -> uint64_t {  // no tcgen05.fence here: TMA wrote the bytes, the tensor core reads them (async proxy both)
        if (!pre) bwait(&w_full[st], ph);
        const int ns = st + 1 == kStackStages ? 0 : st + 1;
        pre = mbar_test_wait(&w_full[ns], ns == 0 ? ph ^ 1 : ph) ? 1u : 0u;
        return dw + static_cast<uint32_t>(st) * kST;
      };
      auto w_release = [&]() {
        if (lead) umma_commit(&w_empty[st]);
        if (++st == kStackStages) { st = 0; ph ^= 1; }
      };
      auto mma4 = [&](uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, bool fresh) {
        if (dbg & 4) return;  // debug: commits only
        if (lead) {
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_bf16(tmem_d, da + 2 * k, db + 2 * k, idesc, (!fresh) || k != 0);
        }
      };
      // head: relu(input_projection(x)) -- A = the cast input in the z buffer                wavenet.py:86-88
      bwait(aux_full, 0);
      tc_fence_after();
      for (int k = 0; k < KBh; ++k) {
        const uint64_t db = w_wait();
        mma4(tm_acc2, dz + k * kZT, db, id64, k == 0);
        w_release();
      }
      if (lead) umma_commit(acc2_full);
      for (int l = 0; l < L; ++l) {
        const uint32_t lp = l & 1;
        const int dil = 1 << (l % p.cycle);
        // GEMM-1: dilated conv, N = 64 gate + 64 filter columns of this CTA's channels         wavenet.py:36-38
        for (int j = 0; j < KB; ++j) {  // centre tap: row offset = the halo
          if (j == 0) {
            bwait(x_loc, lp);
            tc_fence_after();
            mark(l, 0);
          } else {
            bwait(&x_rem[j - 1], lp);
            if (j == KB - 1) mark(l, 1);
          }
          const uint64_t db = w_wait();
          mma4(tm_acc1, dx + ((r + j) % KB) * kXT + kStackHalo * 8, db, id128, j == 0);
          w_release();
        }
        if (has_nb) bwait(halo_full, lp);
        for (int j = 0; j < KB; ++j) {
          const uint64_t da = dx + ((r + j) % KB) * kXT + kStackHalo * 8;
#pragma unroll
          for (int sgn = -1; sgn <= 1; sgn += 2) {
            const uint64_t db = w_wait();
            if (j == KB - 1 && sgn == 1) mark(l, 2);
            mma4(tm_acc1, da + sgn * dil * 8, db, id128, false);
            w_release();
          }
        }
        if (lead) umma_commit(acc1_full);
        // GEMM-2: residual half (fresh accumulator) and skip half (accumulates over all layers)  wavenet.py:44-48,96
        for (int j = 0; j < KB; ++j) {
          if (j == 0) {
            bwait(z_loc, lp);
            tc_fence_after();
            mark(l, 3);
          } else {
            bwait(&z_rem[j - 1], lp);
            if (j == KB - 1) mark(l, 4);
          }
          const uint64_t db = w_wait();
          if (j == KB - 1) mark(l, 5);
          const uint64_t da = dz + ((r + j) % KB) * kZT;
          if (l + 1 < L) mma4(tm_acc2, da, db, id64, j == 0);
          mma4(tm_skip, da, db + (8192 >> 4), id64, l == 0 && j == 0);
          w_release();
        }
        if (lead) umma_commit(acc2_full);
      }
      // relu(skip_projection(sum / sqrt(L))): s1 travels in the x~ buffer (free since the last GEMM-1; the z buffer may still
      // be read by a slower peer's last GEMM-2)                                                  wavenet.py:96-98
      for (int j = 0; j < KB; ++j) {
        if (j == 0) { bwait(x_loc, L & 1); tc_fence_after(); }
        else bwait(&x_rem[j - 1], L & 1);
        const uint64_t db = w_wait();
        mma4(tm_acc2, dx + ((r + j) % KB) * kXT + kStackHalo * 8, db, id64, j == 0);
        w_release();
      }
      if (lead) umma_commit(acc2_full);
      // output_projection: s2 travels in the z buffer                                            wavenet.py:99
      for (int j = 0; j < KB; ++j) {
        if (j == 0) { bwait(z_loc, L & 1); tc_fence_after(); }
        else bwait(&z_rem[j - 1], L & 1);
        const uint64_t db = w_wait();
        mma4(tm_acc1, dz + ((r + j) % KB) * kZT, db, idout, j == 0);
        w_release();
      }
      if (lead) umma_commit(acc1_full);
    }
  } else if (warp == 2) {
    produce(2, elect_one());
  } else if (warp >= 4) {
    // -------------------------------------------------------------- epilogue
    const int quad = warp & 3;
    const int half = (warp - 4) >> 2;
    const int row = quad * 32 + lane;       // frame inside the tile = TMEM lane
    const int t = t0 + row;
    const bool valid = t < Tb;
    const size_t grow = static_cast<size_t>(b) * p.T + (valid ? t : Tb - 1);  // global frame row (clamped for loads)
    const uint32_t lane_addr = tmem_base + (static_cast<uint32_t>(quad * 32) << 16);
    const int c0 = 64 * r + 32 * half;      // first of this thread's 32 channels
    const uint32_t xrow_a = smem_u32(xs) + r * kStackXTile;   // this CTA's k-block of x~
    const uint32_t zrow_a = smem_u32(zs) + r * kStackZTile;
    const bool leader_thread = threadIdx.x == 128;  // warp 4, lane 0
    uint32_t ph1 = 0, ph2 = 0;

    // writes this thread's 32 bf16 channels (16 packed pairs) of its row into k-block tile `tile_a` at row `trow`
    auto put_row = [&](uint32_t tile_a, int trow, const uint32_t (&v)[16]) {
#pragma unroll
      for (int i = 0; i < 4; ++i) st_shared_v4(tile_a + sw128_off(trow, 4 * half + i), v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
    };
    // edge frames of x~ -> flagged packets in the global edge buffer: chunk j (8 channels) of row er = 32 bytes
    //   {ch0-1, flag, ch2-3, flag | ch4-5, flag, ch6-7, flag}; every 8-byte half validates itself
    auto put_edges = [&](int seq, const uint32_t (&v)[16]) {
      if (!has_nb) return;
      const bool top = quad == 0 && lane < kStackHalo && has_left, bot = quad == 3 && lane >= 32 - kStackHalo && has_right;
      if (top || bot) {
        const uint32_t flag = p.flag_base + seq + 1;
        const int side = top ? 0 : 1;
        const int er = top ? lane : lane - (32 - kStackHalo);
        uint8_t* dst = p.edges + ((static_cast<size_t>(seq & 1) * p.n_tiles * 2 + tile * 2 + side) * KB + r) * kStackEdgeBytes +
                       (er * 8 + 4 * half) * 32;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          st_volatile_v4(dst + 32 * i, v[4 * i], flag, v[4 * i + 1], flag);
          st_volatile_v4(dst + 32 * i + 16, v[4 * i + 2], flag, v[4 * i + 3], flag);
        }
      }
    };
    // halo rows of x~ for layer `seq` <- the neighbours' packets.  Runs on all 256 epilogue threads right after they published
    // x~ (they would otherwise idle until the gate accumulator is ready); the previous layer's GEMM-1 -- the last reader of the
    // halo rows -- finished before the epilogue got here (it waited on acc1_full).  One chunk = two 16-byte loads; all of a
    // thread's loads are in flight together, chunks whose flags are not there yet are re-read.
    auto get_halo = [&](int seq) {
      if (!has_nb) return;
      const uint32_t flag = p.flag_base + seq + 1;
      const int te = threadIdx.x - 128;
      constexpr int kPerSide = KB * kStackHalo * 8;          // chunks per side
      constexpr int kMine = (2 * kPerSide + 255) / 256;      // chunks per thread over both sides
      const uint8_t* eb = p.edges + static_cast<size_t>(seq & 1) * p.n_tiles * 2 * KB * kStackEdgeBytes;
      const uint8_t* src[kMine];
      uint32_t dst[kMine];
      uint4 a[kMine], bq[kMine];
#pragma unroll
      for (int i = 0; i < kMine; ++i) {
        const int c = te + 256 * i;
        const int side = c / kPerSide, cc = c % kPerSide;    // side 0: my rows [0, 16) <- the left tile's LAST 16 frames
        const bool on = side < 2 && (side == 0 ? has_left : has_right);
        const int kb = cc / (kStackHalo * 8), rem = cc % (kStackHalo * 8);
        const int nb_tile = side == 0 ? tile - 1 : tile + 1;
        src[i] = on ? eb + (static_cast<size_t>(nb_tile * 2 + (1 - side)) * KB + kb) * kStackEdgeBytes + rem * 32 : nullptr;
        dst[i] = smem_u32(xs) + kb * kStackXTile + sw128_off((side == 0 ? 0 : 128 + kStackHalo) + rem / 8, rem % 8);
        if (on) { a[i] = ld_volatile_v4(src[i]); bq[i] = ld_volatile_v4(src[i] + 16); }
      }
#pragma unroll
      for (int i = 0; i < kMine; ++i) {
        if (src[i] == nullptr) continue;
        uint32_t spins = 0;
        while (a[i].y != flag || a[i].w != flag || bq[i].y != flag || bq[i].w != flag) {
          if (++spins > (1u << 24)) __trap();
          a[i] = ld_volatile_v4(src[i]);
          bq[i] = ld_volatile_v4(src[i] + 16);
        }
        st_shared_v4(dst[i], a[i].x, a[i].z, bq[i].x, bq[i].z);
      }
      fence_proxy_async_smem();
      epi_bar_sync();
      if (leader_thread) mbar_arrive(halo_full);
    };
    // all-gather of this CTA's slice: local arrival + one bulk copy per peer, peers in the order they will consume
    // (the KB-1 copies are issued by lane 0 of KB-1 different warps: issuing one costs a thread ~285 cycles)
    // Copy j goes to the peer that consumes this block at ITS stage j + 1, i.e. completes that peer's bar_rem[j]; the same lane arms
    // this CTA's own bar_rem[j] for the phase being published (`arm`; phase 0 was armed at start-up) -- its previous phase is over:
    // the MMAs that read that block finished before the accumulator this epilogue just consumed was committed.
    auto publish = [&](uint32_t tile_a, uint32_t body_off, uint64_t* bar_loc, uint64_t* bar_rem, bool arm) {
      fence_proxy_async_smem();
      tc_fence_before();
      epi_bar_sync();
      if (elect_one()) {
        if (warp == 5) mbar_arrive(bar_loc);
        if (warp == 4) {  // ONE lane sends all peer copies, in the order the peers consume them: the first lands after 1/(KB-1) of the time
          for (int j = 0; j < KB - 1; ++j) {
            if (arm) {
              if (dbg & 8) mbar_arrive(&bar_rem[j]);
              else mbar_arrive_expect_tx(&bar_rem[j], kRemBytes);
            }
            if (!(dbg & 8)) {
              const uint32_t peer = static_cast<uint32_t>((r + KB - 1 - j) % KB);
              bulk_copy_to_peer(tile_a + body_off, tile_a + body_off, kStackZTile, smem_u32(&bar_rem[j]), peer);
            }
          }
        }
      }
    };

    // ---- head input: x fp32 [B][T][FM] -> bf16 operand tiles in the z buffer (every CTA keeps its own copy)
    grid_dep_wait();
    if (half < KBh) {
      const float* xr = p.x + grow * p.FM + half * 64;
      const uint32_t ta = smem_u32(zs) + half * kStackZTile;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float f[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const int c = half * 64 + j * 8 + e;
          f[e] = (valid && c < p.FM) ? __ldg(xr + j * 8 + e) : 0.0f;
        }
        st_shared_v4(ta + sw128_off(row, j), pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
      }
    }
    fence_proxy_async_smem();
    epi_bar_sync();
    if (leader_thread) mbar_arrive(aux_full);

    // ---- head epilogue: x~_0 = relu(acc + b) + d_0                                              wavenet.py:86-88,36
    {
      float bb[32], dd[32];
      ld_f32x32(p.bhead + c0, bb);
      ld_f32x32(p.dtab + static_cast<size_t>(b) * p.vec_bstride + c0, dd);
      bwait(acc2_full, ph2);
      ph2 ^= 1;
      tc_fence_after();
      uint32_t a[32];
      tmem_ld_32x32(lane_addr + 128 + 32 * half, a);
      tmem_ld_wait();
      uint32_t o[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        const float v0 = valid ? fmaxf(__uint_as_float(a[2 * i]) + bb[2 * i], 0.0f) + dd[2 * i] : 0.0f;
        const float v1 = valid ? fmaxf(__uint_as_float(a[2 * i + 1]) + bb[2 * i + 1], 0.0f) + dd[2 * i + 1] : 0.0f;
        o[i] = pack_bf16(v0, v1);
      }
      put_edges(0, o);
      put_row(xrow_a, row + kStackHalo, o);
      publish(xrow_a, kStackHalo * 128u, x_loc, x_rem, false);
      get_halo(0);
    }

    for (int l = 0; l < L; ++l) {
      // ---- gate: z = sigmoid(g) * tanh(f), packed-half math as in layer_kernel.cuh                wavenet.py:40-42
      {
        const uint16_t* add = p.cproj + (static_cast<size_t>(l) * p.B * p.T + grow) * (2 * C) + 128 * r + 32 * half;
        uint4 bg[4], bf[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          bg[i] = __ldg(reinterpret_cast<const uint4*>(add) + i);
          bf[i] = __ldg(reinterpret_cast<const uint4*>(add + 64) + i);
        }
        bwait(acc1_full, ph1);
        ph1 ^= 1;
        tc_fence_after();
        if (leader_thread) mark(l, 6);
        constexpr uint32_t kHalf2 = 0x38003800u;  // (0.5h, 0.5h)
        uint32_t g[32], f[32];
        tmem_ld_32x32(lane_addr + 32 * half, g);
        tmem_ld_32x32(lane_addr + 64 + 32 * half, f);
        tmem_ld_wait();
        uint32_t zp[16];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const uint32_t bgw[4] = {bg[i].x, bg[i].y, bg[i].z, bg[i].w};
          const uint32_t bfw[4] = {bf[i].x, bf[i].y, bf[i].z, bf[i].w};
#pragma unroll
          for (int h = 0; h < 4; ++h) {
            const int q = 4 * i + h;
            const uint32_t gq = fma_f16x2(pack_f16(__uint_as_float(g[2 * q]), __uint_as_float(g[2 * q + 1])), kHalf2, bgw[h]);
            const uint32_t fq = add_f16x2(pack_f16(__uint_as_float(f[2 * q]), __uint_as_float(f[2 * q + 1])), bfw[h]);
            const uint32_t sg = fma_f16x2(tanh_f16x2(gq), kHalf2, kHalf2);
            zp[q] = f16x2_to_bf16x2(mul_f16x2(sg, tanh_f16x2(fq)));
          }
        }
        put_row(zrow_a, row, zp);
        publish(zrow_a, 0u, z_loc, z_rem, l > 0);
        if (leader_thread) mark(l, 7);
      }
      // ---- residual: x~' = (x~ + R) / sqrt(2) + rvec_l  (not after the last layer: wavenet.py:92-94 discards it)
      if (l + 1 < L) {
        float rr[32];
        ld_f32x32(p.rtab + static_cast<size_t>(b) * p.vec_bstride + static_cast<size_t>(l) * C + c0, rr);
        uint4 xq[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) xq[i] = ld_shared_v4(xrow_a + sw128_off(row + kStackHalo, 4 * half + i));
        bwait(acc2_full, ph2);
        ph2 ^= 1;
        tc_fence_after();
        if (leader_thread) mark(l, 8);
        uint32_t a[32];
        tmem_ld_32x32(lane_addr + 128 + 32 * half, a);
        tmem_ld_wait();
        uint32_t o[16];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const uint32_t w[4] = {xq[i].x, xq[i].y, xq[i].z, xq[i].w};
#pragma unroll
          for (int h = 0; h < 4; ++h) {
            const int q = 8 * i + 2 * h;
            const float v0 = fmaf(__uint_as_float(a[q]) + __uint_as_float(w[h] << 16), 0.70710678118654752f, rr[q]);
            const float v1 = fmaf(__uint_as_float(a[q + 1]) + __uint_as_float(w[h] & 0xffff0000u), 0.70710678118654752f, rr[q + 1]);
            o[4 * i + h] = valid ? pack_bf16(v0, v1) : 0u;
          }
        }
        put_edges(l + 1, o);  // first: the neighbour tile's wait for these packets is the longer path
        put_row(xrow_a, row + kStackHalo, o);
        publish(xrow_a, kStackHalo * 128u, x_loc, x_rem, true);
        if (leader_thread) mark(l, 9);
        get_halo(l + 1);
      }
    }
    // ---- skip sum (TMEM, all layers) -> s1 = (sum + b) / sqrt(L)                                   wavenet.py:96
    {
      float bb[32];
      ld_f32x32(p.bsk + c0, bb);
      bwait(acc2_full, ph2);
      ph2 ^= 1;
      tc_fence_after();
      uint32_t a[32];
      tmem_ld_32x32(lane_addr + 192 + 32 * half, a);
      tmem_ld_wait();
      uint32_t o[16];
#pragma unroll
      for (int i = 0; i < 16; ++i)
        o[i] = pack_bf16((__uint_as_float(a[2 * i]) + bb[2 * i]) * p.inv_sqrt_L, (__uint_as_float(a[2 * i + 1]) + bb[2 * i + 1]) * p.inv_sqrt_L);
      put_row(xrow_a, row + kStackHalo, o);
      publish(xrow_a, kStackHalo * 128u, x_loc, x_rem, true);
    }
    // ---- s2 = relu(skip_projection(s1) + b)                                                        wavenet.py:97-98
    {
      float bb[32];
      ld_f32x32(p.bsp + c0, bb);
      bwait(acc2_full, ph2);
      ph2 ^= 1;
      tc_fence_after();
      uint32_t a[32];
      tmem_ld_32x32(lane_addr + 128 + 32 * half, a);
      tmem_ld_wait();
      uint32_t o[16];
#pragma unroll
      for (int i = 0; i < 16; ++i)
        o[i] = pack_bf16(fmaxf(__uint_as_float(a[2 * i]) + bb[2 * i], 0.0f), fmaxf(__uint_as_float(a[2 * i + 1]) + bb[2 * i + 1], 0.0f));
      put_row(zrow_a, row, o);
      publish(zrow_a, 0u, z_loc, z_rem, true);
    }
    // ---- output projection -> fp32 [B][T][FM]: this CTA's nout columns, 16 per warp half             wavenet.py:99
    {
      const int ncols = p.nout / 2;  // 8 or 16 columns per thread
      const int n0 = p.nout * r + ncols * half;
      const size_t at0 = grow * p.FM + n0;
      // sampler update (if fused): the other operands of up to 3 terms are fetched BEFORE the accumulator wait
      constexpr int kPre = 3;
      float pre[kPre][16];
      const bool prefetch = p.upd_n > 0 && p.upd_n <= kPre && valid;
      if (prefetch) {
#pragma unroll
        for (int k = 0; k < kPre; ++k)
#pragma unroll
          for (int i = 0; i < 16; ++i)
            pre[k][i] = (k < p.upd_n && p.upd_src[k] != p.out && i < ncols && n0 + i < p.FM) ? p.upd_src[k][at0 + i] : 0.0f;
      }
      bwait(acc1_full, ph1);
      ph1 ^= 1;
      tc_fence_after();
      uint32_t a[32];
      tmem_ld_32x32(lane_addr + ncols * half, a);  // reads 32 columns; only the first `ncols` are this thread's
      tmem_ld_wait();
      if (valid) {
        float* orow = p.out + grow * p.FM;
#pragma unroll
        for (int i = 0; i < 16; ++i)
          if (i < ncols && n0 + i < p.FM) {
            const float e = __uint_as_float(a[i]) + __ldg(p.bout + n0 + i);
            orow[n0 + i] = e;
            if (p.upd_n > 0) {  // the sampler's update of this element: lincomb_kernel's fma chain (plain loads: dst may alias a source)
              float acc = 0.0f;
              if (prefetch) {
#pragma unroll
                for (int k = 0; k < kPre; ++k)
                  if (k < p.upd_n) acc = fmaf(p.upd_coef[k], p.upd_src[k] == p.out ? e : pre[k][i], acc);
              } else {
                for (int k = 0; k < p.upd_n; ++k) acc = fmaf(p.upd_coef[k], p.upd_src[k] == p.out ? e : p.upd_src[k][at0 + i], acc);
              }
              p.upd_dst[at0 + i] = acc;
            }
          }
      }
    }
  }
  tc_fence_before();
  cluster_sync_all();  // no CTA leaves while peers may still write its shared memory / read its barriers
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<256>(tmem_base);
  }
}

}  // namespace dsb
