// bm_kernels.cu -- FastSpacedBM for sm_100a: SAD block-matching search + per-axis histogram vote.
//
// Semantics restated from src/FastSpacedBMMethod.cl (OptFlow_C1_D0 :4-84, Histogram_C1_D0 :86-169)
// and the launch geometry of src/FastSpacedBMMethod_OCL.cpp:81-93:
//   block (bx,by): origin x0 = bx*(B+step), y0 = by*(B+step);
//   SAD(sx,sy) = sum_{i,j<B} |curr[y0+R+j][x0+R+i] - prev[y0+j+sy][x0+i+sx]|, sx,sy in [0,2R];
//   argmin = first minimum scanning sy outer / sx inner; shift = (sx-R, sy-R);
//   noise gate: SAD(R,R) - min <= 0.2 R^2 -> (0,0);  vote = per-axis mode, ties -> lowest shift.
//
// Design: one CTA per (block, pair).  The B x B current block and the (B+2R)^2 search window are
// staged in shared memory once; each thread owns one search row sy and four adjacent sx and
// streams over the block with 32-bit words: one broadcast LDS of the current word, one LDS of the
// next window word, three PRMT to build the shifted views, four VABSDIFF4-accumulate (__vsadu4)
// -- 16 pixel-candidates per 9 instructions.  The window pitch is chosen congruent to the number of
// sx-groups modulo 32 so that a warp's window loads hit 32 different banks.  Integer work only:
// results are bit-exact by construction.
#include <cuda.h>
#include <cuda_runtime.h>
#include <limits.h>
#include <stdlib.h>
#include <string.h>

#include "mof_internal.h"

namespace mof {

// ---- TMA staging (cp.async.bulk.tensor + mbarrier; SASS: UTMALDG) ----------------------------------
// The B x B current block and the (B+2R)^2 search window are boxes of a 3-D tensor
// [pair][row][column] of bytes: one elected thread asks the TMA unit for both, the other threads only
// wait on the mbarrier -- no per-byte address arithmetic, no register round trip, and out-of-image
// bytes arrive as zeros.  Needs 16-byte aligned base / row / pair strides and boxes of at most 256 x 256
// (cuTensorMapEncodeTiled's limits); anything else takes the byte loop below.
struct BmTma {
  CUtensorMap curr, prev;
};

__device__ __forceinline__ void mbar_init(uint64_t* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(bar)), "r"(count) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, unsigned parity) {
  const uint32_t a = (uint32_t)__cvta_generic_to_shared(bar);
  unsigned done;
  do {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(done) : "r"(a), "r"(parity) : "memory");
  } while (!done);
}
__device__ __forceinline__ void tma_load_3d(void* dst, const CUtensorMap* map, int x, int y, int z, uint64_t* bar) {
  asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
               ::"r"((uint32_t)__cvta_generic_to_shared(dst)), "l"(map), "r"(x), "r"(y), "r"(z),
                 "r"((uint32_t)__cvta_generic_to_shared(bar)) : "memory");
}

struct BmSmemLayout {
  int pitchC;   // bytes per row of the current block (multiple of 4, >= B + 4)
  int pitchW;   // words per row of the window
  int landC, landW;   // TMA only: bytes per row of the boxes as they land (multiples of 16)
  int offW, offSad, offBar, total;
};
constexpr int BM_KC = 8, BM_KW = 16;   // TMA re-pitch: words a thread carries for the block / the window

// Window rows are read by lane (sy, group of four sx) at word sy*pitchW + group + i: conflict-free only
// when pitchW == number of groups (mod 32), which is odd for the default geometry -- never a multiple of
// 16 bytes; and the TMA unit wants the innermost box coordinate on a 16-byte boundary (measured on this
// B200: an unaligned x raises "illegal instruction", tools/ubench/tma_probe.cu), while blocks start at
// x0 + R.  So both boxes start at the 16-byte boundary below their first pixel, land densely, and are
// re-pitched IN PLACE: every thread takes its words into registers (funnel shift for the sub-word offset),
// one block barrier, every thread stores them at the final pitch -- 23 LDS/STS per thread at B = 120,
// R = 21 against 58 byte loads from global memory in the byte loop.
static __host__ __device__ inline BmSmemLayout bm_layout(int B, int R, bool tma = false, bool wide = false) {
  BmSmemLayout L;
  const int D = 2 * R + 1, G4 = (D + 3) / 4, WW = B + 2 * R;
  L.pitchC = ((B + 3) / 4) * 4 + 4;
  int need = (WW + 3) / 4 + 2;                  // words touched per window row (incl. look-ahead)
  int pw = need + ((G4 - need) % 32 + 32) % 32;  // smallest pw >= need with pw == G4 (mod 32)
  if (wide) {
    // row-wide items (bm_sad_wide_kernel): both arrays are read with 16-byte vector loads by lanes that hold
    // consecutive search rows -- rows on 16-byte boundaries, window pitch == 12 (mod 32) words so that the eight
    // lanes of a quarter warp hit eight different bank quads
    L.pitchC = ((B + 15) / 16) * 16;
    const int needw = ((B / 4 + G4 + 3) / 4) * 4;   // words k = 0 .. B/4 + G4 - 1 (+ look-ahead), whole vectors
    pw = needw + ((12 - needw) % 32 + 32) % 32;
  }
  L.pitchW = pw;
  L.landC = ((B + 15 + 15) / 16) * 16;
  L.landW = ((WW + 15 + 15) / 16) * 16;
  const int szC = tma ? (L.landC > L.pitchC ? L.landC : L.pitchC) * B + 16 : L.pitchC * B;
  const int szW = tma ? (L.landW > pw * 4 ? L.landW : pw * 4) * WW + 16 : pw * 4 * WW;
  L.offW = ((szC + 127) / 128) * 128;
  L.offSad = ((L.offW + szW + 15) / 16) * 16;
  L.offBar = L.offSad + D * D * 4;
  L.offBar = ((L.offBar + 7) / 8) * 8;
  L.total = tma ? L.offBar + 16 : L.offSad + D * D * 4;
  return L;
}

// sum of absolute byte differences added to acc in ONE instruction (VABSDIFF4.U8.ACC); written as
// PTX because `__vsadu4(a, b) + acc` compiles to a zero-accumulator VABSDIFF4 plus an IADD3
__device__ __forceinline__ unsigned sad4(unsigned a, unsigned b, unsigned acc) {
  unsigned r;
  asm("vabsdiff4.u32.u32.u32.add %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(acc));
  return r;
}

// ---- staging of one block and its window (shared by both SAD kernels); ends with a block barrier ---------------
template <bool TMA, int KC, int KW>
__device__ __forceinline__ void bm_stage(const BmArgs& a, const BmTma& tm, const BmSmemLayout& L, unsigned char* sm,
                                         int x0, int y0, int pair) {
  const int B = a.B, R = a.R;
  unsigned char* Cs = sm;
  unsigned char* Ws = sm + L.offW;
  const unsigned char* cimg = a.curr + (size_t)pair * a.pair_stride;
  const unsigned char* pimg = a.prev + (size_t)pair * a.pair_stride;
  const int tid = threadIdx.x, nt = blockDim.x;
  const int WW = B + 2 * R;
  const int pitchWb = L.pitchW * 4;
  // ---- stage: current block and search window --------------------------------------------------
  if (TMA) {
    uint64_t* bar = reinterpret_cast<uint64_t*>(sm + L.offBar);
    if (tid == 0) mbar_init(bar, 1);
    __syncthreads();
    const int xc = x0 + R, xca = xc & ~15, xwa = x0 & ~15;   // boxes start on 16-byte boundaries
    if (tid == 0) {
      mbar_expect_tx(bar, (unsigned)(L.landC * B + L.landW * WW));
      tma_load_3d(Cs, &tm.curr, xca, y0 + R, pair + a.pair0, bar);   // B rows of landC bytes
      tma_load_3d(Ws, &tm.prev, xwa, y0, pair + a.pair0, bar);       // WW rows of landW bytes
    }
    mbar_wait(bar, 0);
    // in-place re-pitch: words -> registers, barrier, registers -> final pitch
    unsigned kc[KC], kw[KW];
    const int wC = L.pitchC / 4, wW = L.pitchW;
    const int dc = xc - xca, dw = x0 - xwa;
#pragma unroll
    for (int k = 0; k < KC; k++) {
      const int e = tid + k * nt;
      kc[k] = 0u;
      if (e < B * wC) {
        const int j = e / wC, i = e - j * wC;
        const unsigned* src = reinterpret_cast<const unsigned*>(Cs + j * L.landC) + ((dc >> 2) + i);
        const unsigned lo = ((dc >> 2) + i) * 4 < L.landC ? src[0] : 0u;
        const unsigned hi = ((dc >> 2) + i + 1) * 4 < L.landC ? src[1] : 0u;
        kc[k] = 4 * i < B ? __funnelshift_r(lo, hi, (dc & 3) * 8) : 0u;
      }
    }
#pragma unroll
    for (int k = 0; k < KW; k++) {
      const int e = tid + k * nt;
      kw[k] = 0u;
      if (e < WW * wW) {
        const int j = e / wW, i = e - j * wW;
        const unsigned* src = reinterpret_cast<const unsigned*>(Ws + j * L.landW) + ((dw >> 2) + i);
        const unsigned lo = ((dw >> 2) + i) * 4 < L.landW ? src[0] : 0u;
        const unsigned hi = ((dw >> 2) + i + 1) * 4 < L.landW ? src[1] : 0u;
        kw[k] = __funnelshift_r(lo, hi, (dw & 3) * 8);   // bytes past WW are never part of a kept candidate
      }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < KC; k++) {
      const int e = tid + k * nt;
      if (e < B * wC) reinterpret_cast<unsigned*>(Cs)[e] = kc[k];
    }
#pragma unroll
    for (int k = 0; k < KW; k++) {
      const int e = tid + k * nt;
      if (e < WW * wW) reinterpret_cast<unsigned*>(Ws)[e] = kw[k];
    }
  } else {
    for (int e = tid; e < B * L.pitchC; e += nt) {
      const int j = e / L.pitchC, i = e - j * L.pitchC;
      Cs[e] = (i < B) ? cimg[(size_t)(y0 + R + j) * a.row_stride + x0 + R + i] : (unsigned char)0;
    }
    for (int e = tid; e < WW * pitchWb; e += nt) {
      const int j = e / pitchWb, i = e - j * pitchWb;
      Ws[e] = (i < WW) ? pimg[(size_t)(y0 + j) * a.row_stride + x0 + i] : (unsigned char)0;
    }
  }
  __syncthreads();

}

// ---- first-minimum argmin (warp 0, __reduce_min_sync), noise gate; shared by both SAD kernels -------------------
__device__ __forceinline__ void bm_argmin(const BmArgs& a, const int* sad, int blk, int pair) {
  const int tid = threadIdx.x, R = a.R, D = 2 * R + 1;
  if (tid < 32) {
    int best = INT_MAX, bidx = INT_MAX;
    for (int e = tid; e < D * D; e += 32) {
      const int v = sad[e];
      if (v < best) { best = v; bidx = e; }   // strict '<' keeps the earliest index per lane
    }
    const int mn = __reduce_min_sync(0xffffffffu, best);
    const int first = __reduce_min_sync(0xffffffffu, best == mn ? bidx : INT_MAX);
    if (tid == 0) {
      int my = first / D, mx = first - my * D;
      int fx = mx - R, fy = my - R;
      if (!a.no_gate && (double)(sad[R * D + R] - mn) <= (double)(R * R) * 0.2) { fx = 0; fy = 0; }
      const int nb = a.gx * a.gy;
      int8_t* o = a.flow_xy + (size_t)pair * 2 * nb;
      o[blk] = (int8_t)fx;
      o[nb + blk] = (int8_t)fy;
    }
  }
}

template <bool TMA>
__global__ void __launch_bounds__(1024, 1) bm_sad_kernel(const BmArgs a, const __grid_constant__ BmTma tm) {
  extern __shared__ __align__(128) unsigned char sm[];
  const int B = a.B, R = a.R, D = 2 * R + 1, G4 = (D + 3) / 4;
  const BmSmemLayout L = bm_layout(B, R, TMA);
  unsigned char* Cs = sm;
  unsigned char* Ws = sm + L.offW;
  int* sad = reinterpret_cast<int*>(sm + L.offSad);
  const int blk = blockIdx.x, pair = blockIdx.y;
  const int by = blk / a.gx, bx = blk - by * a.gx;
  const int pitch = B + a.step;
  const int x0 = bx * pitch, y0 = by * pitch;
  const int tid = threadIdx.x, nt = blockDim.x;
  const int pitchWb = L.pitchW * 4;

  bm_stage<TMA, BM_KC, BM_KW>(a, tm, L, sm, x0, y0, pair);

  // ---- SAD: item = (sy, group of four sx) -----------------------------------------------------
  const int nfull = B / 4;
  const bool has_tail = (B & 3) != 0;
  const unsigned tail_mask = has_tail ? (0xffffffffu >> (8 * (4 - (B & 3)))) : 0xffffffffu;
  for (int item = tid; item < D * G4; item += nt) {
    const int sy = item / G4, g = item - sy * G4;
    unsigned acc0 = 0, acc1 = 0, acc2 = 0, acc3 = 0;
    for (int j = 0; j < B; j++) {
      const unsigned* crow = reinterpret_cast<const unsigned*>(Cs + j * L.pitchC);
      const unsigned* wrow = reinterpret_cast<const unsigned*>(Ws + (j + sy) * pitchWb) + g;
      unsigned w0 = wrow[0];
      // full words: 2 LDS + 3 PRMT + 4 VABSDIFF4.ACC per 16 pixel-candidates, no masking
#pragma unroll 5
      for (int i = 0; i < nfull; i++) {
        const unsigned c = crow[i];
        const unsigned w1 = wrow[i + 1];
        acc0 = sad4(c, w0, acc0);
        acc1 = sad4(c, __byte_perm(w0, w1, 0x4321), acc1);
        acc2 = sad4(c, __byte_perm(w0, w1, 0x5432), acc2);
        acc3 = sad4(c, __byte_perm(w0, w1, 0x6543), acc3);
        w0 = w1;
      }
      if (has_tail) {  // block width not a multiple of 4: drop the bytes past B
        const unsigned c = crow[nfull] & tail_mask;
        const unsigned w1 = wrow[nfull + 1];
        acc0 = sad4(c, w0 & tail_mask, acc0);
        acc1 = sad4(c, __byte_perm(w0, w1, 0x4321) & tail_mask, acc1);
        acc2 = sad4(c, __byte_perm(w0, w1, 0x5432) & tail_mask, acc2);
        acc3 = sad4(c, __byte_perm(w0, w1, 0x6543) & tail_mask, acc3);
      }
    }
    const int sx = 4 * g;
    if (sx + 0 < D) sad[sy * D + sx + 0] = (int)acc0;
    if (sx + 1 < D) sad[sy * D + sx + 1] = (int)acc1;
    if (sx + 2 < D) sad[sy * D + sx + 2] = (int)acc2;
    if (sx + 3 < D) sad[sy * D + sx + 3] = (int)acc3;
  }
  __syncthreads();

  bm_argmin(a, sad, blk, pair);
}

// ---- row-wide items (the default geometry: B = 120, 41 <= 2R+1 <= 44) -------------------------------------------
// bm_sad_kernel spends 3 of its 7 integer-pipe instructions per block word on the shifted views of the window
// (PRMT), and every view is rebuilt by the threads of the neighbouring candidate groups: the integer pipe issues at
// half rate on this part, so those PRMTs are 43 % of the kernel's time.  Here a thread owns ONE search row sy with ALL
// of its NG = 11 candidate groups (44 accumulators) over a slab of RS block rows: per block row it keeps the NW = 30
// words of the current block in registers, streams the window row word by word, builds the three views of a window
// word ONCE and feeds them to every (block word i, group t) with i + t = k -- 1320 VABSDIFF4.ACC + 123 PRMT per row
// instead of 1320 + 990.  Everything is unrolled (register indices are compile-time constants); the partial sums of
// the S = B / RS slabs of a search row meet in shared memory (integer atomics: order-free, bit-exact).
#if !defined(MOF_BMW_VIEW_IMAD)
#define MOF_BMW_VIEW_IMAD 0   // (measured: -4 %, profiles/r4e_ab_bm_views.txt)
#endif
// bytes J .. J + 3 of the pair (lo, hi).  PRMT sits on the integer pipe that VABSDIFF4 saturates; the multiply-add
// form hi32(lo * 2^(32 - 8J)) + hi * 2^(32 - 8J) runs on the otherwise idle FMA pipe (IMAD.HI + IMAD.SHL)
template <int J>
__device__ __forceinline__ unsigned view_shift(unsigned lo, unsigned hi, unsigned mul) {
#if MOF_BMW_VIEW_IMAD
  // `mul` = 2^(32 - 8J) in a register ptxas cannot see through (with the constant it emits LEA.HI: integer pipe again)
  unsigned t, r;
  asm("mul.lo.u32 %0, %1, %2;" : "=r"(t) : "r"(hi), "r"(mul));
  asm("mad.hi.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(lo), "r"(mul), "r"(t));
  return r;
#else
  (void)mul;
  return __byte_perm(lo, hi, J == 1 ? 0x4321 : (J == 2 ? 0x5432 : 0x6543));
#endif
}
constexpr int BMW_NW = 30, BMW_NG = 11, BMW_RS = 15, BMW_THREADS = 352;   // 43 x 8 items on 11 warps
constexpr int BMW_KC = 11, BMW_KW = 21;   // re-pitch words per thread: 120 x 32 / 352, 162 x 44 / 352

template <bool TMA>
__global__ void __launch_bounds__(BMW_THREADS, 2) bm_sad_wide_kernel(const BmArgs a, const __grid_constant__ BmTma tm) {
  extern __shared__ __align__(128) unsigned char sm[];
  const int B = a.B, R = a.R, D = 2 * R + 1;
  const BmSmemLayout L = bm_layout(B, R, TMA, true);
  const unsigned char* Cs = sm;
  const unsigned char* Ws = sm + L.offW;
  int* sad = reinterpret_cast<int*>(sm + L.offSad);
  const int blk = blockIdx.x, pair = blockIdx.y;
  const int by = blk / a.gx, bx = blk - by * a.gx;
  const int pitch = B + a.step;
  const int tid = threadIdx.x;
  for (int e = tid; e < D * D; e += BMW_THREADS) sad[e] = 0;   // (the staging ends with a block barrier)
  bm_stage<TMA, BMW_KC, BMW_KW>(a, tm, L, sm, bx * pitch, by * pitch, pair);

  const int S = B / BMW_RS;
  if (tid < D * S) {
    const int slab = tid / D, sy = tid - slab * D;
    unsigned acc[4 * BMW_NG];
#pragma unroll
    for (int i = 0; i < 4 * BMW_NG; i++) acc[i] = 0u;
    const int pitchWb = L.pitchW * 4;
    const int z = a.no_gate & 2;   // always 0, but neither nvcc nor ptxas can tell: the multipliers stay registers
    const unsigned m1 = 1u << (24 + z), m2 = 1u << (16 + z), m3 = 1u << (8 + z);
#pragma unroll 1
    for (int j = slab * BMW_RS; j < (slab + 1) * BMW_RS; j++) {
      const uint4* crow = reinterpret_cast<const uint4*>(Cs + j * L.pitchC);
      const uint4* wrow = reinterpret_cast<const uint4*>(Ws + (j + sy) * pitchWb);
      unsigned c[32];
#pragma unroll
      for (int q = 0; q < 8; q++) {
        const uint4 v = crow[q];
        c[4 * q] = v.x; c[4 * q + 1] = v.y; c[4 * q + 2] = v.z; c[4 * q + 3] = v.w;
      }
      uint4 cur = wrow[0];
#pragma unroll
      for (int q = 0; q < (BMW_NW + BMW_NG + 2) / 4; q++) {   // window words k = 4q .. 4q + 3
        uint4 nxt = make_uint4(0u, 0u, 0u, 0u);
        if (q + 1 < (BMW_NW + BMW_NG + 2) / 4) nxt = wrow[q + 1];   // (the last word's look-ahead feeds sx = 43 only)
        const unsigned w[5] = {cur.x, cur.y, cur.z, cur.w, nxt.x};
#pragma unroll
        for (int kk = 0; kk < 4; kk++) {
          const int k = 4 * q + kk;
          if (k > BMW_NW + BMW_NG - 2) continue;   // no (i, t) left
          const unsigned v0 = w[kk];
          const unsigned v1 = view_shift<1>(w[kk], w[kk + 1], m1);
          const unsigned v2 = view_shift<2>(w[kk], w[kk + 1], m2);
          const unsigned v3 = view_shift<3>(w[kk], w[kk + 1], m3);
#pragma unroll
          for (int t = 0; t < BMW_NG; t++) {
            const int i = k - t;
            if (i < 0 || i >= BMW_NW) continue;
            acc[4 * t + 0] = sad4(c[i], v0, acc[4 * t + 0]);
            acc[4 * t + 1] = sad4(c[i], v1, acc[4 * t + 1]);
            acc[4 * t + 2] = sad4(c[i], v2, acc[4 * t + 2]);
            acc[4 * t + 3] = sad4(c[i], v3, acc[4 * t + 3]);
          }
        }
        cur = nxt;
      }
    }
    int* srow = sad + sy * D;
#pragma unroll
    for (int i = 0; i < 4 * BMW_NG; i++)
      if (i < D) atomicAdd(&srow[i], (int)acc[i]);
  }
  __syncthreads();
  bm_argmin(a, sad, blk, pair);
}

// per-axis histogram vote: one warp per pair
__global__ void __launch_bounds__(32) bm_vote_kernel(const BmArgs a) {
  __shared__ int hx[64], hy[64];
  const int pair = blockIdx.x, lane = threadIdx.x;
  const int R = a.R, D = 2 * R + 1, nb = a.gx * a.gy;
  for (int e = lane; e < 64; e += 32) { hx[e] = 0; hy[e] = 0; }
  __syncwarp();
  const int8_t* f = a.flow_xy + (size_t)pair * 2 * nb;
  for (int e = lane; e < nb; e += 32) {
    atomicAdd(&hx[f[e] + R], 1);
    atomicAdd(&hy[f[nb + e] + R], 1);
  }
  __syncwarp();
  // mode with ties -> lowest bin: maximise count, then minimise bin
  int bcx = -1, bbx = INT_MAX, bcy = -1, bby = INT_MAX;
  for (int e = lane; e < D; e += 32) {
    if (hx[e] > bcx) { bcx = hx[e]; bbx = e; }
    if (hy[e] > bcy) { bcy = hy[e]; bby = e; }
  }
  const int mx = __reduce_max_sync(0xffffffffu, bcx);
  const int ix = __reduce_min_sync(0xffffffffu, bcx == mx ? bbx : INT_MAX);
  const int my = __reduce_max_sync(0xffffffffu, bcy);
  const int iy = __reduce_min_sync(0xffffffffu, bcy == my ? bby : INT_MAX);
  if (lane == 0) {
    a.voted_xy[(size_t)pair * 2 + 0] = (float)(ix - R);
    a.voted_xy[(size_t)pair * 2 + 1] = (float)(iy - R);
  }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn g_encode = nullptr;

int bm_setup(int device) {
  cudaError_t e = cudaSetDevice(device);
  if (e != cudaSuccess) return (int)e;
  e = cudaFuncSetAttribute(bm_sad_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
  if (e != cudaSuccess) return (int)e;
  e = cudaFuncSetAttribute(bm_sad_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
  if (e != cudaSuccess) return (int)e;
  e = cudaFuncSetAttribute(bm_sad_wide_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  if (e != cudaSuccess) return (int)e;
  e = cudaFuncSetAttribute(bm_sad_wide_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  if (e != cudaSuccess) return (int)e;
  if (!g_encode) {   // the driver's tensor-map encoder, through the runtime (no -lcuda)
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      g_encode = reinterpret_cast<EncodeTiledFn>(fn);
    (void)cudaGetLastError();
  }
  return 0;
}

// [pair][row][column] byte tensor with a (box_w x box_h x 1) box; false when TMA cannot describe it
static bool bm_make_map(CUtensorMap* m, const uint8_t* base, int width, int height, int n_pairs, int row_stride,
                        size_t pair_stride, int box_w, int box_h) {
  if (!g_encode) return false;
  if (((size_t)base & 15) || (row_stride & 15) || (pair_stride & 15) || box_w > 256 || box_h > 256 || (box_w & 15))
    return false;
  const cuuint64_t dims[3] = {(cuuint64_t)width, (cuuint64_t)height, (cuuint64_t)n_pairs};
  const cuuint64_t strides[2] = {(cuuint64_t)row_stride, (cuuint64_t)pair_stride};
  const cuuint32_t box[3] = {(cuuint32_t)box_w, (cuuint32_t)box_h, 1u};
  const cuuint32_t estr[3] = {1u, 1u, 1u};
  return g_encode(m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, const_cast<uint8_t*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

int bm_launch(const BmArgs& a, cudaStream_t stream, int* launches) {
  const int D = 2 * a.R + 1, G4 = (D + 3) / 4;
  BmTma tm;
  // the default geometry (120 px blocks, 11 candidate groups) runs the row-wide kernel (MOF_BM_NO_WIDE=1: A/B, tests)
  const bool wide = a.B == 4 * BMW_NW && G4 == BMW_NG && a.B % BMW_RS == 0 && D * (a.B / BMW_RS) <= BMW_THREADS &&
                    !getenv("MOF_BM_NO_WIDE");
  BmSmemLayout L = bm_layout(a.B, a.R, true, wide);
  // TMA staging when the geometry allows it (MOF_BM_NO_TMA=1 forces the byte loop: A/B, tests)
  int threads = ((D * G4 + 31) / 32) * 32;
  if (threads > 1024) threads = 1024;
  if (threads < 64) threads = 64;
  if (wide) threads = BMW_THREADS;
  const int kc_max = wide ? BMW_KC : BM_KC, kw_max = wide ? BMW_KW : BM_KW;
  // TMA staging when the geometry allows it (MOF_BM_NO_TMA=1 forces the byte loop: A/B, tests)
  bool tma = !getenv("MOF_BM_NO_TMA") && L.total <= 227 * 1024 &&
             (a.B * (L.pitchC / 4) + threads - 1) / threads <= kc_max &&
             ((a.B + 2 * a.R) * L.pitchW + threads - 1) / threads <= kw_max &&
             bm_make_map(&tm.curr, a.curr, a.width, a.height, a.n_pairs, a.row_stride, a.pair_stride, L.landC, a.B) &&
             bm_make_map(&tm.prev, a.prev, a.width, a.height, a.n_pairs, a.row_stride, a.pair_stride, L.landW,
                         a.B + 2 * a.R);
  if (!tma) {
    L = bm_layout(a.B, a.R, false, wide);
    memset(&tm, 0, sizeof(tm));
  }
  if (L.total > 227 * 1024) return (int)cudaErrorInvalidValue;
  const int nb = a.gx * a.gy;
  // gridDim.y is limited to 65535: chunk the pairs
  for (int p0 = 0; p0 < a.n_pairs; p0 += 65535) {
    BmArgs b = a;
    const int np = (a.n_pairs - p0 < 65535) ? a.n_pairs - p0 : 65535;
    b.curr = a.curr + (size_t)p0 * a.pair_stride;
    b.prev = a.prev + (size_t)p0 * a.pair_stride;
    b.flow_xy = a.flow_xy + (size_t)p0 * 2 * nb;
    b.voted_xy = a.voted_xy + (size_t)p0 * 2;
    b.n_pairs = np;
    b.pair0 = p0;
    if (wide && tma) bm_sad_wide_kernel<true><<<dim3(nb, np), threads, L.total, stream>>>(b, tm);
    else if (wide) bm_sad_wide_kernel<false><<<dim3(nb, np), threads, L.total, stream>>>(b, tm);
    else if (tma) bm_sad_kernel<true><<<dim3(nb, np), threads, L.total, stream>>>(b, tm);
    else bm_sad_kernel<false><<<dim3(nb, np), threads, L.total, stream>>>(b, tm);
    if (launches) (*launches)++;
  }
  bm_vote_kernel<<<a.n_pairs, 32, 0, stream>>>(a);
  if (launches) (*launches)++;
  return (int)cudaGetLastError();
}

}  // namespace mof
