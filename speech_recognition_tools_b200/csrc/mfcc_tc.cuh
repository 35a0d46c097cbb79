// Kc on the 5th-generation tensor cores: the MFCC path's 513-point DFT as a GEMM (tcgen05.mma + TMA + TMEM).
//
// Reference: src/featgen/computeMfccFeatures.py:111-131 -- spec = abs(fft(frames, nfft/2 + 1)) is a P = 513 POINT
// transform of the 320-sample Hamming frame (no power-of-two FFT exists for it), all P bins meet the filterbank, then
// log10 and the first 13 coefficients of a DCT-II.  csrc/mfcc.cuh evaluates the 257 distinct bins as a direct DFT on
// the FP32 pipe (280 GFLOP per bench batch, 10.8 ms).  Here the same contraction runs on the tensor cores:
//
//   C[F x 512] = A[F x 320] . B[320 x 512]      A = windowed frames / 2^15, B = cos | -sin of bins 1..256
//
// in the 3xTF32 scheme (A = Ah + Al, B = Bh + Bl with Ah, Bh the values cut to 10 mantissa bits -- what kind::tf32
// reads of a 32-bit word -- and C = Ah Bh + Ah Bl + Al Bh accumulated in FP32: the error of an FP32 dot product; one
// TF32 pass alone misses the 2e-4 tolerance of the cepstra, profiles/mel_gemm_probe_r02.json).  Bin 0 is the sum of
// the frame and comes from the preparation kernel.
//
//   mfcc_tc_prep_kernel  one warp per frame: reflect-padded samples (+ alpha * noise) times the window -> Ah, Al
//                        rows (float32) and the frame sum
//   mfcc_tc_kernel       persistent, one CTA per SM, warp roles:
//                          warp 0    TMA producer: per k-block of 32 samples the 128-frame tiles of Ah / Al and the
//                                    256-column tiles of Bh / Bl (128-byte swizzle), 2 stages of 96 KB
//                          warp 1    MMA issuer: one lane issues 3 x tcgen05.mma (M 128, N 256, K 8) per k-step into
//                                    one of two 256-column accumulators in tensor memory; tcgen05.commit frees the
//                                    stage and, after the last k-block, hands the accumulator to the epilogue
//                          warps 2-5 epilogue: thread = frame (TMEM lane); tcgen05.ld 32 columns of re and of im,
//                                    magnitudes, then the folded filterbank for the bands that touch the 32-bin chunk
//                                    (weights broadcast from shared memory, 32 band sums in registers over the two
//                                    halves of the spectrum: unit 0 = bins 1..128, unit 1 = bins 129..256), then bin 0,
//                                    log10 and the DCT-II, all in float32
//                        The accumulator of unit u + 1 fills while the epilogue drains unit u.
//
// Measured on the bench batch (852 138 frames, profiles/ncu_full_r02_mfcc_tc_summary.json): preparation 0.60 ms (2.1 GB
// written, 49 % of HBM) + GEMM kernel 1.14 ms (tensor pipe 75 % busy) = 1.75 ms against 10.8 ms of the FP32-pipe kernel.
// What the first working version (2.8 ms, tensor pipe 38 %) was missing, found with per-role cycle counters
// (-DFDLP_X_TC_PROF): the MMA warp waited for the epilogue half of the time.  (1) The epilogue used FP64 (log10 and the
// DCT in double, as csrc/mfcc.cuh does): DFMA and the tensor cores share a datapath on this part, and behind a stream of
// tcgen05.mma 416 DFMAs per frame took 16 k cycles per tile, 32 double logarithms 30 k -- as long as the tile's MMAs
// (32 k).  Now float log10 and a four-chain compensated float DCT (~2e-5 on c0).  (2) All 256 x 32 filterbank weights per
// frame: 2048 16-byte shared-memory reads; now only the ~100 (32-bin chunk, band) blocks with a non-zero weight, as a
// branch-free loop (the block's sum is added through a select over the 32 register-resident band sums: a warp-uniform
// branch per band serialised load and FMA latencies).  Rejected on measurements: per-band 8-column tcgen05.ld segments
// (latency-bound, 4.9 ms); the tables as a 27 KB kernel parameter in the constant bank (indexed reads slower than shared
// memory, 2.6 ms); an L2 prefetch of the next tile's frames (no change); and a variant without the preparation kernel --
// four more warps building the Ah / Al tiles of a stage directly from the waveform, thread = frame, swizzled 16-byte
// stores, fence.proxy.async -- which was correct and SLOWER (7.5 ms): a frame's 32 samples of a k-block are 64 unaligned
// bytes, 320 bytes from the next frame's, so the builders issue 2-byte loads that touch 32 sectors per warp instruction;
// staging a tile's sample span in shared memory first has no room beside the two 96 KB stages; and four 48 KB stages of
// 16-sample k-blocks with the 64-byte swizzle (-DFDLP_TC_BK=16: correct, 1.74 against 1.66 ms -- smaller boxes, twice the
// barrier traffic).  What bounds the 75 %: bytes INTO the SM.  A k-block brings 96 KB for 12 MMAs = 1600 tensor cycles, i.e.
// 60 bytes per cycle at full rate against an ingest of ~64 per SM; two variants that cut L2 reads but not the bytes
// entering each SM changed nothing: pairs of CTAs sharing every B tile by TMA multicast (-DFDLP_TC_CLUSTER=2: cluster
// launch, multicast loads and multicast tcgen05.commit on the stage barriers: correct, 1.66-1.67 against 1.67-1.71 ms)
// and L2 eviction hints that keep the frames between a tile's two units (-DFDLP_TC_A_HINT: 1.70 ms).  Only a larger
// tile per byte would move it (cta_group::2: each SM loads half of B), or fewer bytes per product.
#pragma once
#include <cuda.h>

#include "common.cuh"

namespace fdlp {
namespace tc {

constexpr int kM = 128;                 // frames per tile
constexpr int kNUnit = 256;             // accumulator columns per unit: 128 bins x (re | im)
constexpr int kUnits = 2;
constexpr int kBinsUnit = 128;
constexpr int kK = 320;                 // samples per frame
#ifndef FDLP_TC_BK
#define FDLP_TC_BK 32
#endif
#ifndef FDLP_TC_CLUSTER
#define FDLP_TC_CLUSTER 1
#endif
constexpr int kCluster = FDLP_TC_CLUSTER;  // 2: pairs of CTAs (consecutive tiles) share every B tile by TMA multicast
constexpr int kBK = FDLP_TC_BK;         // tf32 elements per k-block = one swizzle row: 32 (128-byte swizzle) or 16 (64-byte)
constexpr int kRowBytes = kBK * 4;
constexpr int kKB = kK / kBK;
constexpr int kStages = (kBK == 32) ? 2 : 4;
constexpr int kATile = kM * kRowBytes;  // bytes of one A plane of a stage
constexpr int kBTile = kNUnit * kRowBytes;
constexpr int kStageBytes = 2 * kATile + 2 * kBTile;
constexpr int kJP = 32;                 // bands, padded
constexpr int kThreads = 192;
constexpr int kChunks = kUnits * kBinsUnit / 32;   // 32-bin chunks of the spectrum
constexpr int kMaxBlk = 176;                       // (chunk, band) blocks with a non-zero weight (mel, 30 bands: ~100)
constexpr int kMaxCep = 16;
constexpr size_t kTabBytes = (size_t)kMaxBlk * 32 * 4 + kJP * 4 + (kChunks + 1) * 4 + kMaxBlk + 4 + (size_t)kMaxCep * kJP * 4;
constexpr size_t kSmemBytes = 1024 + (size_t)kStages * kStageBytes + kTabBytes + 128;

struct Args {
  const float* dc;       // [n_frames] frame sums (bin 0)
  // folded filterbank of bins 1..256 in 32-bin chunks: the (chunk, band) blocks with a non-zero weight (a mel band
  // touches 1-3 of the 8 chunks, directly or mirrored), chunk after chunk: blocks blk_off[ch] .. blk_off[ch+1] belong to
  // chunk ch, block e to band blk_band[e], its 32 weights are wblk[e]
  const float* wblk;         // [n_blk][32]
  const uint8_t* blk_band;   // [n_blk]
  const int32_t* blk_off;    // [9]
  int n_blk;
  const float* w0;       // [32] weights of bin 0
  const double* dctb;    // [ncep][32] DCT-II basis, zero beyond J
  float* feats;          // [n_frames, ncep]
  int64_t n_frames;
  int J, ncep, n_tiles;
};

struct PrepArgs {
  const void* wav;
  const int64_t* utt_wav_off;
  const int64_t* utt_frame_off;
  const int32_t* frame_utt;
  int64_t n_frames;
  int flen, shift;
  const float* ham;              // [320] hamming(flen) / 2^15
  float* ah;                     // [n_frames, 320]
  float* al;
  float* dc;                     // [n_frames]
  const int16_t* noise;
  const int64_t* utt_noise_off;
  const double* utt_alpha;
};

template <typename S>
__global__ void __launch_bounds__(256) mfcc_tc_prep_kernel(const PrepArgs a) {
  const int lane = threadIdx.x & 31;
  const int64_t fr = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (fr >= a.n_frames) return;
  const int u = a.frame_utt[fr];
  const int k = (int)(fr - a.utt_frame_off[u]);
  const int64_t base = a.utt_wav_off[u];
  const int64_t T = a.utt_wav_off[u + 1] - base;
  const int ext = (a.flen % 2 == 0) ? a.flen / 2 - 1 : (a.flen - 1) / 2;  // features.py:137-146
  const int64_t q0 = (int64_t)k * a.shift - ext;
  const S* wav = reinterpret_cast<const S*>(a.wav) + base;
  const int16_t* nz = a.noise ? a.noise + a.utt_noise_off[u] : nullptr;
  const float alpha = a.noise ? (float)a.utt_alpha[u] : 0.f;
  float sum = 0.f;
  const bool inside = q0 >= 0 && q0 + kK <= T;  // all but the first and last frames of an utterance: no reflection (a
#pragma unroll                                   // 64-bit modulo per sample otherwise)
  for (int t = 0; t < kK / 32; ++t) {
    const int i = lane + 32 * t;
    const int64_t q = inside ? q0 + i : ((T > 1) ? reflect_index(q0 + i, T) : 0);
    float x = (float)wav[q];
    if (nz) x = fmaf(alpha, (float)nz[q], x);
    const float v = x * __ldg(a.ham + i);
    const float hi = __uint_as_float(__float_as_uint(v) & 0xffffe000u);
    a.ah[fr * kK + i] = hi;
    a.al[fr * kK + i] = v - hi;
    sum += v;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  if (lane == 0) a.dc[fr] = sum;
}

// ---------------------------------------------------------------- PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
#ifdef FDLP_X_TC_PROF
__device__ long long g_tc_prof[8];
#define TC_PROF_ADD(slot, v) atomicAdd((unsigned long long*)&g_tc_prof[slot], (unsigned long long)(v))
#else
#define TC_PROF_ADD(slot, v)
#endif
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  const long long t0 = clock64();
  for (;;) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.b32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    if (ok) return;
    if (clock64() - t0 > 8000000000LL) __trap();  // a lost arrival must not hang the device (seconds, never in a correct run)
  }
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* tm, uint32_t bar, int x, int y) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(dst),
      "l"(reinterpret_cast<uint64_t>(tm)), "r"(bar), "r"(x), "r"(y)
      : "memory");
}
// the same with an L2 eviction-priority hint (policy words as CUTLASS's TMA::CacheHintSm90 encodes them)
constexpr uint64_t kEvictFirst = 0x12F0000000000000ull, kEvictLast = 0x14F0000000000000ull;
__device__ __forceinline__ void tma_load_2d_hint(uint32_t dst, const CUtensorMap* tm, uint32_t bar, int x, int y, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%3, %4}], [%2], %5;" ::"r"(dst),
      "l"(reinterpret_cast<uint64_t>(tm)), "r"(bar), "r"(x), "r"(y), "l"(policy)
      : "memory");
}
__device__ __forceinline__ void tma_load_2d_mc(uint32_t dst, const CUtensorMap* tm, uint32_t bar, int x, int y, uint16_t mask) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1, {%3, %4}], [%2], %5;" ::"r"(dst),
      "l"(reinterpret_cast<uint64_t>(tm)), "r"(bar), "r"(x), "r"(y), "h"(mask)
      : "memory");
}
__device__ __forceinline__ void umma_commit_mc(uint32_t bar, uint16_t mask) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar), "h"(mask)
               : "memory");
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void umma_tf32(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc)
      : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }

// 32 consecutive columns of this thread's TMEM lane
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t* v) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
        "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
        "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// shared-memory matrix descriptor of a K-major tile with 128-byte rows and the 128-byte swizzle (what the TMA box
// {32 floats, rows} writes): 8-row groups 1024 bytes apart, descriptor version 1 (sm_100)
__device__ __forceinline__ uint64_t smem_desc(uint32_t saddr) {
  uint64_t d = (uint64_t)((saddr & 0x3ffffu) >> 4);
  d |= (uint64_t)1 << 16;            // leading byte offset: not used by swizzled K-major layouts
  d |= (uint64_t)((8 * kRowBytes) >> 4) << 32;  // stride byte offset: 8 rows
  d |= (uint64_t)1 << 46;                       // version
  d |= (uint64_t)(kBK == 32 ? 2 : 4) << 61;     // SWIZZLE_128B / SWIZZLE_64B
  return d;
}

// instruction descriptor: D = F32, A = B = TF32, both K-major, N = 256, M = 128
constexpr uint32_t kIdesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(kNUnit >> 3) << 17) | ((uint32_t)(kM >> 4) << 24);

#if FDLP_TC_CLUSTER == 2
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
#else
__global__ void __launch_bounds__(kThreads, 1)
#endif
mfcc_tc_kernel(const __grid_constant__ CUtensorMap tm_ah, const __grid_constant__ CUtensorMap tm_al,
               const __grid_constant__ CUtensorMap tm_bh, const __grid_constant__ CUtensorMap tm_bl, const Args a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  // the epilogue's tables: shared memory, broadcast reads (as a 27 KB kernel parameter in the constant bank the indexed
  // reads were slower: 2.62 against 2.27 ms)
  float* dct_s = reinterpret_cast<float*>(base + (size_t)kStages * kStageBytes);   // [kMaxCep][32]
  float* wblk = dct_s + kMaxCep * kJP;                                            // [kMaxBlk][32]
  float* w0 = wblk + kMaxBlk * 32;                                                // [32]
  int32_t* blk_off = reinterpret_cast<int32_t*>(w0 + kJP);                        // [9] first block of a chunk
  uint8_t* blk_band = reinterpret_cast<uint8_t*>(blk_off + kChunks + 1);          // [kMaxBlk] band of a block
  uint64_t* bars = reinterpret_cast<uint64_t*>(blk_band + kMaxBlk + 4);  // full[S] empty[S] tmem_full[2] tmem_empty[2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * kStages + 4);
  const uint32_t sbase = smem_u32(base);
  const uint32_t bar0 = smem_u32(bars);
  auto full_bar = [&](int s) { return bar0 + 8u * s; };
  auto empty_bar = [&](int s) { return bar0 + 8u * (kStages + s); };
  auto tfull_bar = [&](int u) { return bar0 + 8u * (2 * kStages + u); };
  auto tempty_bar = [&](int u) { return bar0 + 8u * (2 * kStages + 2 + u); };
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  for (int i = threadIdx.x; i < a.n_blk * 32; i += kThreads) wblk[i] = __ldg(a.wblk + i);
  for (int i = threadIdx.x; i < a.n_blk; i += kThreads) blk_band[i] = a.blk_band[i];
  for (int i = threadIdx.x; i < a.ncep * kJP; i += kThreads) dct_s[i] = (float)__ldg(a.dctb + i);
  for (int i = threadIdx.x; i < kJP; i += kThreads) w0[i] = __ldg(a.w0 + i);
  if (threadIdx.x <= kChunks) blk_off[threadIdx.x] = a.blk_off[threadIdx.x];
  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), kCluster); }
    for (int u = 0; u < kUnits; ++u) { mbar_init(tfull_bar(u), 1); mbar_init(tempty_bar(u), 4); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {  // tensor memory: all 512 columns (one CTA per SM)
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512u)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  if (kCluster > 1) cluster_sync_all();  // the peer's barriers exist before anything is multicast to them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int crank = (kCluster > 1) ? (int)(blockIdx.x & 1u) : 0;  // rank in the CTA pair; the pair walks tiles tb, tb + 1
  const int tstart = (int)blockIdx.x - crank;
#ifdef FDLP_X_TC_PROF
  const long long t_start = clock64();
#endif

  if (warp == 0) {
    // ================= TMA producer =================
    if (lane == 0) {
      uint32_t it = 0;
      for (int tb = tstart; tb < a.n_tiles; tb += gridDim.x) {
        const int tile = tb + crank;  // beyond the last tile for the odd CTA of the last pair: zero rows, nothing stored
        for (int u = 0; u < kUnits; ++u)
          for (int kb = 0; kb < kKB; ++kb, ++it) {
            const int s = it % kStages;
            const uint32_t ph = (it / kStages) & 1u;
            { const long long tp_ = clock64(); mbar_wait(empty_bar(s), ph ^ 1u); if (true) TC_PROF_ADD(0, clock64() - tp_); }
            const uint32_t dst = sbase + (uint32_t)s * kStageBytes;
            mbar_expect_tx(full_bar(s), kStageBytes);
#ifdef FDLP_TC_A_HINT   // keep the frames in L2 between the two units of a tile, drop them after the second
            const uint64_t pol = (u == 0) ? kEvictLast : kEvictFirst;
            tma_load_2d_hint(dst, &tm_ah, full_bar(s), kb * kBK, tile * kM, pol);
            tma_load_2d_hint(dst + kATile, &tm_al, full_bar(s), kb * kBK, tile * kM, pol);
#else
            tma_load_2d(dst, &tm_ah, full_bar(s), kb * kBK, tile * kM);
            tma_load_2d(dst + kATile, &tm_al, full_bar(s), kb * kBK, tile * kM);
#endif
            if (kCluster == 1) {
              tma_load_2d(dst + 2 * kATile, &tm_bh, full_bar(s), kb * kBK, u * kNUnit);
              tma_load_2d(dst + 2 * kATile + kBTile, &tm_bl, full_bar(s), kb * kBK, u * kNUnit);
            } else {  // this CTA fetches its half of the rows of Bh and Bl for both CTAs of the pair
              const int half = kNUnit / 2;
              tma_load_2d_mc(dst + 2 * kATile + crank * (kBTile / 2), &tm_bh, full_bar(s), kb * kBK, u * kNUnit + crank * half, 3);
              tma_load_2d_mc(dst + 2 * kATile + kBTile + crank * (kBTile / 2), &tm_bl, full_bar(s), kb * kBK,
                             u * kNUnit + crank * half, 3);
            }
          }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    uint32_t it = 0, tile_it = 0;
    for (int tb = tstart; tb < a.n_tiles; tb += gridDim.x, ++tile_it)
      for (int u = 0; u < kUnits; ++u) {
        { const long long tp_ = clock64(); mbar_wait(tempty_bar(u), (tile_it & 1u) ^ 1u); if (lane == 0) TC_PROF_ADD(1, clock64() - tp_); }  // the epilogue has drained this accumulator
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)u * kNUnit;
        for (int kb = 0; kb < kKB; ++kb, ++it) {
          const int s = it % kStages;
          { const long long tp_ = clock64(); mbar_wait(full_bar(s), (it / kStages) & 1u); if (lane == 0) TC_PROF_ADD(2, clock64() - tp_); }
          tc_fence_after();
          if (lane == 0) {
            const uint32_t st = sbase + (uint32_t)s * kStageBytes;
            const uint64_t dah = smem_desc(st), dal = smem_desc(st + kATile);
            const uint64_t dbh = smem_desc(st + 2 * kATile), dbl = smem_desc(st + 2 * kATile + kBTile);
#pragma unroll
            for (int ks = 0; ks < kBK / 8; ++ks) {
              const uint64_t o = (uint64_t)(ks * 32 >> 4);  // 8 tf32 = 32 bytes along the swizzled row
              umma_tf32(d_tmem, dah + o, dbh + o, kIdesc, (kb | ks) ? 1u : 0u);
              umma_tf32(d_tmem, dah + o, dbl + o, kIdesc, 1u);
              umma_tf32(d_tmem, dal + o, dbh + o, kIdesc, 1u);
            }
            if (kCluster == 1) umma_commit(empty_bar(s));   // the stage is free once these MMAs have read it
            else umma_commit_mc(empty_bar(s), 3);           // ... for both producers of the pair (each writes into both)
            if (kb == kKB - 1) umma_commit(tfull_bar(u));   // and the accumulator is complete
          }
          __syncwarp();
        }
      }
  } else {
    // ================= epilogue: thread = frame =================
    const int quarter = warp & 3;  // the TMEM lanes a warp may read: 32 * (warp id % 4)
    const int row = quarter * 32 + lane;
    uint32_t tile_it = 0;
    for (int tb = tstart; tb < a.n_tiles; tb += gridDim.x, ++tile_it) {
      const int tile = tb + crank;
      float band[kJP];
#pragma unroll
      for (int j = 0; j < kJP; ++j) band[j] = 0.f;
      for (int u = 0; u < kUnits; ++u) {
        { const long long tp_ = clock64(); mbar_wait(tfull_bar(u), tile_it & 1u); if (threadIdx.x == 64) TC_PROF_ADD(3, clock64() - tp_); }
        tc_fence_after();
        const uint32_t t0 = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)u * kNUnit;
        for (int c = 0; c < kBinsUnit / 32; ++c) {
          uint32_t re[32], im[32];
#ifdef FDLP_X_TC_PROF
          const long long tl_ = clock64();
#endif
          tmem_ld32(t0 + c * 32, re);
          tmem_ld32(t0 + kBinsUnit + c * 32, im);
          tmem_ld_wait();
#ifdef FDLP_X_TC_PROF
          const long long tm_ = clock64();
#endif
          float mag[32];
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            const float xr = __uint_as_float(re[i]), xi = __uint_as_float(im[i]);
            const float p2 = fmaf(xr, xr, xi * xi);
            mag[i] = p2 > 0.f ? p2 * rsqrtf(p2) : 0.f;  // 2 ulp: the log10 of a sum of ~30 of these is compared to 2e-4
          }
          // only the (chunk, band) blocks with a non-zero weight, ~100 of the 256 of a mel filterbank, as one runtime loop
          // (a chunk's blocks are consecutive in wblk; the band a block belongs to is a table entry, and its sum is added
          // through a select over the 32 register-resident band sums: branch-free, so the loads of the next block overlap
          // the FMAs of this one -- a warp-uniform branch per band serialised load and FMA latencies, 4.3 k cycles per chunk)
          const int ch = u * (kBinsUnit / 32) + c;
          const int e0 = blk_off[ch], e1 = blk_off[ch + 1];
#pragma unroll 2
          for (int e = e0; e < e1; ++e) {
            const float4* wrow = reinterpret_cast<const float4*>(wblk + (size_t)e * 32);
            const int je = blk_band[e];
            float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
            for (int q = 0; q < 8; ++q) {
              const float4 w = wrow[q];
              a0 = fmaf(w.x, mag[4 * q + 0], a0);
              a1 = fmaf(w.y, mag[4 * q + 1], a1);
              a2 = fmaf(w.z, mag[4 * q + 2], a2);
              a3 = fmaf(w.w, mag[4 * q + 3], a3);
            }
            const float part = (a0 + a1) + (a2 + a3);
#pragma unroll
            for (int j = 0; j < kJP; ++j) band[j] += (j == je) ? part : 0.f;
          }
#ifdef FDLP_X_TC_PROF
          if (threadIdx.x == 64) TC_PROF_ADD(6, clock64() - tm_);
#endif
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(tempty_bar(u));
      }
      const int64_t fr = (int64_t)tile * kM + row;
#ifdef FDLP_X_TC_PROF
      const long long tf_ = clock64();
#endif
      if (fr < a.n_frames) {
        const float dc = fabsf(__ldg(a.dc + fr));
        // No FP64 here: DFMA and the tensor cores share a datapath on this part, and behind a stream of
        // tcgen05.mma the 416 DFMAs of the double-precision DCT took 16 k cycles per tile, 32 double logarithms another
        // 30 k.  float log10 of the float band energy, float products and a compensated (Kahan) sum: ~2e-5 on c0.
        float le[kJP];
#pragma unroll
        for (int j = 0; j < kJP; ++j) le[j] = (j < a.J) ? log10f(fmaf(w0[j], dc, band[j])) : 0.f;
#ifdef FDLP_X_TC_PROF
        if (threadIdx.x == 64) TC_PROF_ADD(5, clock64() - tf_);
#endif
        for (int c = 0; c < a.ncep; ++c) {
          const float* brow = dct_s + (size_t)c * kJP;  // broadcast reads
          float sq[4] = {0.f, 0.f, 0.f, 0.f}, cq[4] = {0.f, 0.f, 0.f, 0.f};  // four compensated chains
#pragma unroll
          for (int j = 0; j < kJP; ++j) {
            const float y = brow[j] * le[j] - cq[j & 3];
            const float t = sq[j & 3] + y;
            cq[j & 3] = (t - sq[j & 3]) - y;
            sq[j & 3] = t;
          }
          const float s = ((sq[0] + sq[1]) + (sq[2] + sq[3])) - ((cq[0] + cq[1]) + (cq[2] + cq[3]));
          a.feats[fr * a.ncep + c] = s;
        }
      }
#ifdef FDLP_X_TC_PROF
      if (threadIdx.x == 64) TC_PROF_ADD(7, clock64() - tf_);
#endif
    }
  }
#ifdef FDLP_X_TC_PROF
  if (threadIdx.x == 0) TC_PROF_ADD(4, clock64() - t_start);
#endif
  tc_fence_before();
  __syncthreads();
  if (kCluster > 1) cluster_sync_all();  // no multicast or remote arrival is still on its way to this CTA
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
  }
}

}  // namespace tc
}  // namespace fdlp
