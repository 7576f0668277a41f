// pg_kernels.cu -- phase correlation for section sizes outside the prime-factor plan (see pg_generic.cuh):
// any even N = 2^a 3^b 5^c up to 768, which covers the reference constructor's own fallback N := frameSize
// (src/FftMethod.cpp:1709-1716: one 480 px section at config/default.yaml's frame size) and the 30 / 80 /
// 96 / 160 px sections users configure.  Same boundary and epilogue as pc_pfa_kernel (PcArgs in, shifts
// out); a persistent CTA of 16 warps works on one section at a time:
//   rows:     warp w transforms rows w, w+16, ... of z = curr + i prev in its shared-memory line buffer and
//             writes them to the CTA's N x N complex scratch (L2 / HBM, coalesced);
//   columns:  tiles of 16 columns are scattered into the 16 line buffers (128 B per scratch row and tile),
//             one warp per column, gathered back;
//   cross-power on bin pairs (k, -k), inverse columns, inverse rows with a running first maximum, the real
//   surface stored over the row's own scratch, 5x5 FP64 centroid by warp 0.
// Stream pushes keep the previous FRAME of a stream (not its spectrum) in the context's cache: a push is a
// pairs launch whose prev pointers go through the stream ids, followed by pg_cache_kernel.
#include <cuda_runtime.h>

#include "mof_internal.h"
#include "pg_generic.cuh"

namespace mof {

using namespace pg;

// warps per CTA = columns per column tile: 16 where two CTAs fit an SM's shared memory (N <= 400), 8 above
// (three CTAs of 8 warps instead of one of 16: +13 % at N = 480)
static int pg_tile(int M) { return M > 400 ? 8 : 16; }
static size_t pg_smem_bytes(int M, int tile) { return ((size_t)M + (size_t)tile * pg_pitch(M)) * sizeof(c32); }

template <bool INV>
__device__ __forceinline__ void pg_line_fft(c32* buf, const c32* tw, const PgPlan& pl, int lane) {
  for (int k = 0; k < pl.nr; k++) {
    pg_line_pass<INV>(buf, tw, pl, k, lane);
    __syncwarp();
  }
}

template <bool INV, int PG_TILE>
__device__ __forceinline__ void pg_columns(c32* Z, c32* s_buf, const c32* s_tw, const PgPlan& pl, int tid) {
  constexpr int PG_THREADS = PG_TILE * 32;
  const int M = pl.M, pitch = pg_pitch(M), res = pg_result_half(pl) * pg_half(M);
  const int warp = tid >> 5, lane = tid & 31;
  for (int x0 = 0; x0 < M; x0 += PG_TILE) {
    const int ncol = M - x0 < PG_TILE ? M - x0 : PG_TILE;
    for (int idx = tid; idx < M * PG_TILE; idx += PG_THREADS) {
      const int row = idx / PG_TILE, col = idx - row * PG_TILE;
      if (col < ncol) s_buf[col * pitch + row] = Z[(size_t)row * M + x0 + col];
    }
    __syncthreads();
    if (warp < ncol) pg_line_fft<INV>(s_buf + warp * pitch, s_tw, pl, lane);
    __syncthreads();
    for (int idx = tid; idx < M * PG_TILE; idx += PG_THREADS) {
      const int row = idx / PG_TILE, col = idx - row * PG_TILE;
      if (col < ncol) Z[(size_t)row * M + x0 + col] = s_buf[col * pitch + res + row];
    }
    __syncthreads();
  }
}

// cache: previous frames of the streams (F x F bytes each, tight) or nullptr; out_fs: sections between the
// results of consecutive frames (S, or T * S when the call writes frame t of a [n][T][S] array)
template <int PG_TILE>
__global__ void __launch_bounds__(PG_TILE * 32, PG_TILE == 16 ? 2 : 3)
pg_kernel(const __grid_constant__ PcArgs a, const __grid_constant__ PgPlan pl, const c32* __restrict__ g_tw,
          c32* __restrict__ scratch, const uint8_t* __restrict__ cache, int out_fs) {
  constexpr int PG_THREADS = PG_TILE * 32;
  extern __shared__ __align__(16) unsigned char pg_smem[];
  const int M = pl.M, pitch = pg_pitch(M), res = pg_result_half(pl) * pg_half(M);
  c32* s_tw = reinterpret_cast<c32*>(pg_smem);
  c32* s_buf = s_tw + M;
  __shared__ float s_best[PG_TILE];
  __shared__ int s_bidx[PG_TILE];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int j = tid; j < M; j += PG_THREADS) s_tw[j] = g_tw[j];
  c32* Z = scratch + (size_t)blockIdx.x * M * M;
  c32* buf = s_buf + warp * pitch;
  const int F = a.sq * M;
  __syncthreads();

  for (int sec = blockIdx.x; sec < a.n_sections; sec += gridDim.x) {
    const int f = sec / a.S, s_ = sec - f * a.S;
    const int sj = s_ / a.sq, si = s_ - sj * a.sq;
    const uint8_t* curr = a.curr + (size_t)f * a.frame_stride + (size_t)(sj * M) * a.row_stride + si * M;
    const uint8_t* prev;
    int prs = a.row_stride;
    if (a.mode == PC_PAIRS) {
      prev = a.prev + (size_t)f * a.frame_stride + (size_t)(sj * M) * a.row_stride + si * M;
    } else if (a.inline_flags[f] & 1u) {   // first call of the stream: imPrev := imCurr (src/FftMethod.cpp:1791-1793)
      prev = curr;
    } else {
      prev = cache + (size_t)a.inline_ids[f] * F * F + (size_t)(sj * M) * F + si * M;
      prs = F;
    }
    // forward rows of z = curr + i prev
    for (int y = warp; y < M; y += PG_TILE) {
      for (int x = lane; x < M; x += 32) {
        float c = (float)curr[(size_t)y * a.row_stride + x], p = (float)prev[(size_t)y * prs + x];
        if (a.win) {
          const float w = a.win[y * M + x];
          c *= w;
          p *= w;
        }
        buf[x] = mk(c, p);
      }
      __syncwarp();
      pg_line_fft<false>(buf, s_tw, pl, lane);
      for (int x = lane; x < M; x += 32) Z[(size_t)y * M + x] = buf[res + x];
      __syncwarp();
    }
    __syncthreads();
    pg_columns<false, PG_TILE>(Z, s_buf, s_tw, pl, tid);
    // cross-power: one visit per bin pair (k, -k)
    const int hw = M / 2 + 1;
    for (int idx = tid; idx < M * hw; idx += PG_THREADS) {
      const int ky = idx / hw, kx = idx - ky * hw;
      const int my = ky ? M - ky : 0, mx = kx ? M - kx : 0;
      if (mx == kx && my < ky) continue;
      const bool self = mx == kx && my == ky;
      const c32 c = pg_cross(Z[(size_t)ky * M + kx], Z[(size_t)my * M + mx], self);
      Z[(size_t)ky * M + kx] = c;
      if (!self) Z[(size_t)my * M + mx] = mk(c.x, -c.y);
    }
    __syncthreads();
    pg_columns<true, PG_TILE>(Z, s_buf, s_tw, pl, tid);
    // inverse rows: real surface over the row's own scratch, running first maximum
    float best = -INFINITY;
    int bidx = 0x7fffffff;
    for (int y = warp; y < M; y += PG_TILE) {
      for (int x = lane; x < M; x += 32) buf[x] = Z[(size_t)y * M + x];
      __syncwarp();
      pg_line_fft<true>(buf, s_tw, pl, lane);
      float* srow = reinterpret_cast<float*>(Z + (size_t)y * M);
      for (int x = lane; x < M; x += 32) {
        const float v = buf[res + x].x;
        srow[x] = v;
        pg_max_update(v, pg_shifted_index(y, x, M), best, bidx);
      }
      __syncwarp();
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) {
      const float ob = __shfl_xor_sync(0xffffffffu, best, o);
      const int oi = __shfl_xor_sync(0xffffffffu, bidx, o);
      pg_max_update(ob, oi, best, bidx);
    }
    if (lane == 0) { s_best[warp] = best; s_bidx[warp] = bidx; }
    __syncthreads();
    if (warp == 0) {
      best = lane < PG_TILE ? s_best[lane] : -INFINITY;
      bidx = lane < PG_TILE ? s_bidx[lane] : 0x7fffffff;
#pragma unroll
      for (int o = 16; o; o >>= 1) {
        const float ob = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bidx, o);
        pg_max_update(ob, oi, best, bidx);
      }
      double cx = 0.0, cy = 0.0, sum = 0.0;
      if (bidx != 0x7fffffff && lane < 25) {
        const int ys = bidx / M + lane / 5 - 2, xs = bidx % M + lane % 5 - 2;
        if (ys >= 0 && ys < M && xs >= 0 && xs < M) {
          const float v = reinterpret_cast<const float*>(Z + (size_t)pg_unshift(ys, M) * M)[pg_unshift(xs, M)];
          pg_acc(ys, xs, v, cx, cy, sum);
        }
      }
#pragma unroll
      for (int o = 16; o; o >>= 1) {
        cx += __shfl_xor_sync(0xffffffffu, cx, o);
        cy += __shfl_xor_sync(0xffffffffu, cy, o);
        sum += __shfl_xor_sync(0xffffffffu, sum, o);
      }
      if (lane == 0) {
        double xy[2];
        if (bidx == 0x7fffffff) {   // no maximum: the surface is all NaN
          const bool one = a.post == PC_POST_SCALEROT && !a.sr_reject_nan;
          xy[0] = one ? 1.0 : (double)NAN;
          xy[1] = one ? 0.0 : (double)NAN;
        }
        else if (a.post == PC_POST_SCALEROT) pg_finish_scalerot(cx, cy, sum, M, a.sr_m, a.sr_reject_nan, xy);
        else pg_finish(cx, cy, sum, M, a.max_px_speed_sq, xy);
        double* o = a.out + ((size_t)f * out_fs + s_) * 2;
        o[0] = xy[0];
        o[1] = xy[1];
      }
    }
    __syncthreads();   // the scratch and s_best are reused by the next section
  }
}

// frames of one push -> the streams' previous-frame cache (after pg_kernel has read it)
__global__ void pg_cache_kernel(const __grid_constant__ PcArgs a, uint8_t* __restrict__ cache, int F) {
  const int f = blockIdx.x;
  const uint8_t* src = a.curr + (size_t)f * a.frame_stride;
  uint8_t* dst = cache + (size_t)a.inline_ids[f] * F * F;
  if (((((size_t)src) | (size_t)a.row_stride | (size_t)F) & 3) == 0) {
    const int wpr = F >> 2;
    for (int i = threadIdx.x + blockIdx.y * blockDim.x; i < F * wpr; i += blockDim.x * gridDim.y) {
      const int y = i / wpr, w = i - y * wpr;
      reinterpret_cast<uint32_t*>(dst)[i] = reinterpret_cast<const uint32_t*>(src + (size_t)y * a.row_stride)[w];
    }
  } else {
    for (int i = threadIdx.x + blockIdx.y * blockDim.x; i < F * F; i += blockDim.x * gridDim.y) {
      const int y = i / F, x = i - y * F;
      dst[i] = src[(size_t)y * a.row_stride + x];
    }
  }
}

bool pg_supported(int N) {
  PgPlan pl;
  return pg_plan(N, &pl);
}

size_t pg_scratch_bytes_per_cta(int N) { return (size_t)N * N * sizeof(c32); }

// resident CTAs for section size N on `device` (also opts the kernel into its shared memory); <= 0 on error
int pg_resident_ctas(int N, int device) {
  PgPlan pl;
  if (!pg_plan(N, &pl)) return 0;
  if (cudaSetDevice(device) != cudaSuccess) return 0;
  int sms = 0, per = 0;
  if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device) != cudaSuccess) return 0;
  const int tile = pg_tile(N);
  const size_t smem = pg_smem_bytes(N, tile);
  const int smax = (int)pg_smem_bytes(PG_MAX_N, 8);
  cudaError_t e = tile == 16 ? cudaFuncSetAttribute(pg_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pg_smem_bytes(400, 16))
                             : cudaFuncSetAttribute(pg_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, smax);
  if (e != cudaSuccess) return 0;
  e = tile == 16 ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per, pg_kernel<16>, 512, smem)
                 : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per, pg_kernel<8>, 256, smem);
  if (e != cudaSuccess) return 0;
  if (per > 4) per = 4;
  return sms * per;
}

// twiddles W_N^j = exp(-2 pi i j / N), j < N, in double precision on the host (2 * N floats)
void pg_twiddles(int N, float* out) {
  for (int j = 0; j < N; j++) {
    const double ang = -2.0 * 3.14159265358979323846 * (double)j / (double)N;
    out[2 * j] = (float)cos(ang);
    out[2 * j + 1] = (float)sin(ang);
  }
}

int pg_launch(int N, const PcArgs& args, const float* d_tw, void* scratch, const uint8_t* cache, int out_fs, int grid,
              cudaStream_t stream) {
  PgPlan pl;
  if (!pg_plan(N, &pl)) return (int)cudaErrorInvalidValue;
  if (args.mode != PC_PAIRS && args.mode != PC_STREAM) return (int)cudaErrorInvalidValue;
  if (grid < 1) return 0;
  if (pg_tile(N) == 16)
    pg_kernel<16><<<grid, 512, pg_smem_bytes(N, 16), stream>>>(args, pl, reinterpret_cast<const c32*>(d_tw),
                                                               reinterpret_cast<c32*>(scratch), cache, out_fs);
  else
    pg_kernel<8><<<grid, 256, pg_smem_bytes(N, 8), stream>>>(args, pl, reinterpret_cast<const c32*>(d_tw),
                                                             reinterpret_cast<c32*>(scratch), cache, out_fs);
  return (int)cudaGetLastError();
}

int pg_cache_update(const PcArgs& args, int n_frames, uint8_t* cache, int F, cudaStream_t stream) {
  if (n_frames < 1) return 0;
  pg_cache_kernel<<<dim3(n_frames, 8), 256, 0, stream>>>(args, cache, F);
  return (int)cudaGetLastError();
}

}  // namespace mof
