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
#include <cuda_runtime.h>
#include <limits.h>

#include "mof_internal.h"

namespace mof {

struct BmSmemLayout {
  int pitchC;   // bytes per row of the current block (multiple of 4, >= B + 4)
  int pitchW;   // words per row of the window
  int offW, offSad, total;
};

static __host__ __device__ inline BmSmemLayout bm_layout(int B, int R) {
  BmSmemLayout L;
  const int D = 2 * R + 1, G4 = (D + 3) / 4;
  L.pitchC = ((B + 3) / 4) * 4 + 4;
  int need = (B + 2 * R + 3) / 4 + 2;          // words touched per window row (incl. look-ahead)
  int pw = need + ((G4 - need) % 32 + 32) % 32;  // smallest pw >= need with pw == G4 (mod 32)
  L.pitchW = pw;
  L.offW = L.pitchC * B;
  L.offSad = L.offW + pw * 4 * (B + 2 * R);
  L.total = L.offSad + D * D * 4;
  return L;
}

// sum of absolute byte differences added to acc in ONE instruction (VABSDIFF4.U8.ACC); written as
// PTX because `__vsadu4(a, b) + acc` compiles to a zero-accumulator VABSDIFF4 plus an IADD3
__device__ __forceinline__ unsigned sad4(unsigned a, unsigned b, unsigned acc) {
  unsigned r;
  asm("vabsdiff4.u32.u32.u32.add %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(acc));
  return r;
}

__global__ void __launch_bounds__(1024, 1) bm_sad_kernel(const BmArgs a) {
  extern __shared__ __align__(16) unsigned char sm[];
  const int B = a.B, R = a.R, D = 2 * R + 1, G4 = (D + 3) / 4;
  const BmSmemLayout L = bm_layout(B, R);
  unsigned char* Cs = sm;
  unsigned char* Ws = sm + L.offW;
  int* sad = reinterpret_cast<int*>(sm + L.offSad);
  const int blk = blockIdx.x, pair = blockIdx.y;
  const int by = blk / a.gx, bx = blk - by * a.gx;
  const int pitch = B + a.step;
  const int x0 = bx * pitch, y0 = by * pitch;
  const unsigned char* cimg = a.curr + (size_t)pair * a.pair_stride;
  const unsigned char* pimg = a.prev + (size_t)pair * a.pair_stride;
  const int tid = threadIdx.x, nt = blockDim.x;
  const int WW = B + 2 * R;
  const int pitchWb = L.pitchW * 4;

  // ---- stage: current block (zero padded to the word pitch) and search window -------------------
  for (int e = tid; e < B * L.pitchC; e += nt) {
    const int j = e / L.pitchC, i = e - j * L.pitchC;
    Cs[e] = (i < B) ? cimg[(size_t)(y0 + R + j) * a.row_stride + x0 + R + i] : (unsigned char)0;
  }
  for (int e = tid; e < WW * pitchWb; e += nt) {
    const int j = e / pitchWb, i = e - j * pitchWb;
    Ws[e] = (i < WW) ? pimg[(size_t)(y0 + j) * a.row_stride + x0 + i] : (unsigned char)0;
  }
  __syncthreads();

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

  // ---- first-minimum argmin (warp 0, __reduce_min_sync), noise gate ------------------------------
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

int bm_setup(int device) {
  cudaError_t e = cudaSetDevice(device);
  if (e != cudaSuccess) return (int)e;
  e = cudaFuncSetAttribute(bm_sad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
  return (int)e;
}

int bm_launch(const BmArgs& a, cudaStream_t stream, int* launches) {
  const int D = 2 * a.R + 1, G4 = (D + 3) / 4;
  const BmSmemLayout L = bm_layout(a.B, a.R);
  if (L.total > 227 * 1024) return (int)cudaErrorInvalidValue;
  int threads = ((D * G4 + 31) / 32) * 32;
  if (threads > 1024) threads = 1024;
  if (threads < 64) threads = 64;
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
    bm_sad_kernel<<<dim3(nb, np), threads, L.total, stream>>>(b);
    if (launches) (*launches)++;
  }
  bm_vote_kernel<<<a.n_pairs, 32, 0, stream>>>(a);
  if (launches) (*launches)++;
  return (int)cudaGetLastError();
}

}  // namespace mof
