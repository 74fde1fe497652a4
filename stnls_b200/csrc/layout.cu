// Layout kernels: reference planar video (B,HD,T,F,H,W) / (B,T,HD*F,H,W)  <->  channels-last
// (B*HD, Tloc, H, W, Fp).  This is the one copy the reference also makes in `shape_vids`
// (lib/stnls/search/utils.py:178-190, 'b t (HD c) h w -> b HD t c h w' + .contiguous()); here it
// also pads channels to a multiple of 4 so every kernel moves float4 channel vectors, and (for the
// aggregation backward) fuses the division by the patch-overlap counts (lib/stnls/agg/gather_add.py:133-134).
#include "common.cuh"

namespace {

struct LayoutParams {
  float* cl;
  float* planar;
  int HD, t0, Tloc, F, Fp, H, W;
  int64_t sB, sHD, sT, sC;
  const int* counts_h;
  const int* counts_w;
  int accumulate;
  int Tpitch, tcl0;   // the channels-last video holds Tpitch frames per (b, head); the Tloc frames start at frame tcl0
                      // (a rank packs its OWN frames straight into the owned + halo buffer its peers read)
  int64_t nz;   // B*HD*Tloc slabs
};

// Several videos in ONE launch (q, k, v of a step; the three gradients on the way back), or one: blockIdx.z walks the
// slabs (b, head, frame) of all jobs; a block whose (x, y) lies outside a smaller job's frame skips it.
// A block moves a strip of 32 pixels x R rows x Fp channels through a padded shared tile [R][Fp][33]: on the planar side
// every warp reads / writes 32 consecutive pixels of one channel and keeps R rows in flight per thread, on the
// channels-last side the strip's rows are contiguous runs of 32*Fp floats moved as float4.  (Round 1 moved ONE row per
// block -- 1 KB at 8 channels -- and ran at 0.55 TB/s on the README shape: block scheduling, not bandwidth.)
constexpr int LAYOUT_MAX_JOBS = 4;
struct LayoutJobs {
  LayoutParams job[LAYOUT_MAX_JOBS];
  int64_t zend[LAYOUT_MAX_JOBS];   // cumulative slab counts
  int njobs;
};

template <bool PACK, int R>
__global__ void __launch_bounds__(256) layout_kernel(const __grid_constant__ LayoutJobs J) {
  extern __shared__ __align__(16) float tile[];  // [R][max Fp][33]
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int w0 = blockIdx.x * 32, h0 = blockIdx.y * R;
  const int64_t total = J.zend[J.njobs - 1];
  for (int64_t zz = blockIdx.z; zz < total; zz += gridDim.z) {
    int j = 0;
    while (zz >= J.zend[j]) j++;
    const LayoutParams& P = J.job[j];
    if (w0 >= P.W || h0 >= P.H) continue;   // block-uniform
    const int64_t z = zz - (j ? J.zend[j - 1] : 0);
    const int tl = (int)(z % P.Tloc), n = (int)(z / P.Tloc);
    const int b = n / P.HD, hd = n % P.HD;
    const int Fp = P.Fp;
    float* pl = P.planar + (int64_t)b * P.sB + (int64_t)hd * P.sHD + (int64_t)(P.t0 + tl) * P.sT + (int64_t)h0 * P.W + w0 + tx;
    float* cl = P.cl + ((((int64_t)n * P.Tpitch + P.tcl0 + tl) * P.H + h0) * P.W + w0) * Fp;
    const int wn = min(32, P.W - w0), hr = min(R, P.H - h0);
    const int per = wn * (Fp >> 2);          // float4s of one strip row on the channels-last side
    const int64_t clrow = (int64_t)P.W * Fp;
    if (PACK) {
      const bool div = P.counts_h != nullptr;
      float inv[R];
#pragma unroll
      for (int r = 0; r < R; r++)
        inv[r] = (div && r < hr && tx < wn) ? (float)(P.counts_h[h0 + r] * P.counts_w[w0 + tx]) : 1.f;
      for (int c = ty; c < Fp; c += 8) {
        float v[R];
#pragma unroll
        for (int r = 0; r < R; r++)
          v[r] = (c < P.F && r < hr && tx < wn) ? __ldg(pl + (int64_t)c * P.sC + (int64_t)r * P.W) : 0.f;
#pragma unroll
        for (int r = 0; r < R; r++) tile[(r * Fp + c) * 33 + tx] = div ? v[r] / inv[r] : v[r];
      }
      __syncthreads();
      for (int i = threadIdx.x; i < hr * per; i += 256) {
        const int r = i / per, rem = i - r * per;
        const int wl = (rem << 2) / Fp, c = (rem << 2) - wl * Fp;
        const float* tp = tile + (r * Fp + c) * 33 + wl;
        *reinterpret_cast<float4*>(cl + r * clrow + ((int64_t)rem << 2)) = make_float4(tp[0], tp[33], tp[66], tp[99]);
      }
    } else {
      for (int i = threadIdx.x; i < hr * per; i += 256) {
        const int r = i / per, rem = i - r * per;
        const int wl = (rem << 2) / Fp, c = (rem << 2) - wl * Fp;
        const float4 x = __ldg(reinterpret_cast<const float4*>(cl + r * clrow + ((int64_t)rem << 2)));
        float* tp = tile + (r * Fp + c) * 33 + wl;
        tp[0] = x.x; tp[33] = x.y; tp[66] = x.z; tp[99] = x.w;
      }
      __syncthreads();
      for (int c = ty; c < P.F; c += 8) {
#pragma unroll
        for (int r = 0; r < R; r++) {
          if (r < hr && tx < wn) {
            float* d = pl + (int64_t)c * P.sC + (int64_t)r * P.W;
            const float v = tile[(r * Fp + c) * 33 + tx];
            *d = P.accumulate ? (*d + v) : v;
          }
        }
      }
    }
    __syncthreads();   // the tile is reused by the next slab
  }
}

__global__ void add_inplace_kernel(float* dst, const float* src, int64_t n) {
  int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i + 3 < n) {
    float4 a = *reinterpret_cast<float4*>(dst + i);
    float4 b = *reinterpret_cast<const float4*>(src + i);
    a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
    *reinterpret_cast<float4*>(dst + i) = a;
  } else {
    for (; i < n; i++) dst[i] += src[i];
  }
}

__global__ void add_inplace_scalar_kernel(float* dst, const float* src, int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dst[i] += src[i];
}

// up to 8 contiguous segment copies (or accumulations) in ONE launch: the halo exchange stages / pulls the frames of
// several tensors and several neighbours at once (stnls_b200/parallel.py PeerHalo).  Sources may be peer-GPU pointers.
struct SegJobs {
  float* dst[8];
  const float* src[8];
  int64_t n4[8];      // float4 elements of one row (segments are 16-byte multiples, 16-byte aligned)
  int rows[8];        // rows of the segment (1: a plain contiguous copy) ...
  int64_t dpitch[8], spitch[8];   // ... and the distance between rows, in floats, on either side
  int nseg, accumulate;
};
__global__ void __launch_bounds__(256) seg_copy_kernel(const __grid_constant__ SegJobs J) {
  const int sgm = blockIdx.y;
  const int64_t n4 = J.n4[sgm], total = n4 * J.rows[sgm];
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const float* sbase = J.src[sgm];
  float* dbase = J.dst[sgm];
  // rows x n4 float4s as one index space; four independent 16-byte loads in flight per thread (peer loads over NVLink
  // have microseconds of latency: parallelism, not per-thread speed, is what fills the link)
  for (int64_t i0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i0 < total; i0 += 4 * stride) {
    float4 v[4];
    int64_t off_d[4];
    bool ok[4];
#pragma unroll
    for (int u = 0; u < 4; u++) {
      const int64_t i = i0 + u * stride;
      ok[u] = i < total;
      const int64_t r = ok[u] ? i / n4 : 0, c = ok[u] ? i - r * n4 : 0;
      off_d[u] = r * J.dpitch[sgm] + c * 4;
      if (ok[u]) v[u] = *reinterpret_cast<const float4*>(sbase + r * J.spitch[sgm] + c * 4);
    }
#pragma unroll
    for (int u = 0; u < 4; u++) {
      if (!ok[u]) continue;
      // accumulate: two neighbours may return gradient for the same frame -> vector RED, not a read-modify-write
      if (J.accumulate) red_add4(dbase + off_d[u], v[u]);
      else *reinterpret_cast<float4*>(dbase + off_d[u]) = v[u];
    }
  }
}

template <bool PACK, int R>
int launch_layout_r(const LayoutJobs& J, dim3 grid, size_t smem, cudaStream_t st, const char* who) {
  if (smem > 48 * 1024) {   // more than 372 channels per head: opt in to the large carve-out
    cudaError_t e = cudaFuncSetAttribute(layout_kernel<PACK, R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    STNLS_REQUIRE(e == cudaSuccess, "%s: cannot reserve %zu bytes of shared memory", who, smem);
  }
  layout_kernel<PACK, R><<<grid, 256, smem, st>>>(J);
  STNLS_CHECK_LAUNCH(who);
  return STNLS_OK;
}

// J.job[0..njobs) filled except nz / zend; B per job
int launch_layout_jobs(bool pack, LayoutJobs& J, const int* B, cudaStream_t st, const char* who) {
  int64_t z = 0;
  int maxW = 0, maxH = 0, maxFp = 0;
  for (int j = 0; j < J.njobs; j++) {
    LayoutParams& P = J.job[j];
    STNLS_REQUIRE(P.cl && P.planar, "%s: null argument", who);
    STNLS_REQUIRE(P.Fp % 4 == 0 && P.Fp >= P.F && P.F > 0, "%s: Fp must be a multiple of 4 and >= F", who);
    STNLS_REQUIRE(P.HD > 0 && B[j] >= 0 && P.Tloc >= 0 && P.H >= 0 && P.W >= 0, "%s: bad shape", who);
    STNLS_REQUIRE((reinterpret_cast<uintptr_t>(P.cl) & 15) == 0, "%s: the channels-last video must be 16-byte aligned", who);
    P.nz = (P.H == 0 || P.W == 0) ? 0 : (int64_t)B[j] * P.HD * P.Tloc;
    z += P.nz;
    J.zend[j] = z;
    maxW = P.W > maxW ? P.W : maxW; maxH = P.H > maxH ? P.H : maxH; maxFp = P.Fp > maxFp ? P.Fp : maxFp;
  }
  if (z == 0) return STNLS_OK;
  const int R = maxFp <= 16 ? 8 : maxFp <= 32 ? 4 : maxFp <= 64 ? 2 : 1;
  const size_t smem = (size_t)R * maxFp * 33 * sizeof(float);
  STNLS_REQUIRE(smem <= 227 * 1024, "%s: too many channels per head (%d)", who, maxFp);
  STNLS_REQUIRE((maxH + R - 1) / R <= 65535, "%s: H too large", who);
  dim3 grid((maxW + 31) / 32, (maxH + R - 1) / R, (unsigned)(z < 65535 ? z : 65535));
  if (pack) {
    if (R == 8) return launch_layout_r<true, 8>(J, grid, smem, st, who);
    if (R == 4) return launch_layout_r<true, 4>(J, grid, smem, st, who);
    if (R == 2) return launch_layout_r<true, 2>(J, grid, smem, st, who);
    return launch_layout_r<true, 1>(J, grid, smem, st, who);
  }
  if (R == 8) return launch_layout_r<false, 8>(J, grid, smem, st, who);
  if (R == 4) return launch_layout_r<false, 4>(J, grid, smem, st, who);
  if (R == 2) return launch_layout_r<false, 2>(J, grid, smem, st, who);
  return launch_layout_r<false, 1>(J, grid, smem, st, who);
}

int launch_layout(bool pack, LayoutParams& P, int B, cudaStream_t st, const char* who) {
  LayoutJobs J;
  J.job[0] = P;
  J.njobs = 1;
  return launch_layout_jobs(pack, J, &B, st, who);
}

}  // namespace

extern "C" int stnls_pack_cl(float* dst_cl, const float* src, int B, int HD, int t0, int Tloc, int F, int Fp, int H,
                             int W, int64_t sB, int64_t sHD, int64_t sT, int64_t sC, const int* counts_h,
                             const int* counts_w, stnls_stream_t stream) {
  LayoutParams P;
  P.cl = dst_cl; P.planar = const_cast<float*>(src); P.HD = HD; P.t0 = t0; P.Tloc = Tloc; P.F = F; P.Fp = Fp;
  P.H = H; P.W = W; P.sB = sB; P.sHD = sHD; P.sT = sT; P.sC = sC; P.counts_h = counts_h; P.counts_w = counts_w;
  P.accumulate = 0; P.Tpitch = Tloc; P.tcl0 = 0;
  STNLS_REQUIRE((counts_h == nullptr) == (counts_w == nullptr), "pack_cl: counts_h and counts_w go together");
  return launch_layout(true, P, B, (cudaStream_t)stream, "pack_cl");
}

extern "C" int stnls_unpack_cl(float* dst, const float* src_cl, int B, int HD, int t0, int Tloc, int F, int Fp, int H,
                               int W, int64_t sB, int64_t sHD, int64_t sT, int64_t sC, int accumulate,
                               stnls_stream_t stream) {
  LayoutParams P;
  P.cl = const_cast<float*>(src_cl); P.planar = dst; P.HD = HD; P.t0 = t0; P.Tloc = Tloc; P.F = F; P.Fp = Fp;
  P.H = H; P.W = W; P.sB = sB; P.sHD = sHD; P.sT = sT; P.sC = sC; P.counts_h = nullptr; P.counts_w = nullptr;
  P.accumulate = accumulate; P.Tpitch = Tloc; P.tcl0 = 0;
  return launch_layout(false, P, B, (cudaStream_t)stream, "unpack_cl");
}

// host side of the multi-job launch; job arrays are HOST arrays of njobs entries, strides 4 per job (sB,sHD,sT,sC)
static int launch_layout_multi(bool pack, int njobs, float* const* cl, float* const* planar, const int* B, const int* HD,
                               const int* t0, const int* Tloc, const int* F, const int* H, const int* W,
                               const int64_t* strides, const int* const* counts_h, const int* const* counts_w,
                               const int* accumulate, const int* cl_tpitch, const int* cl_t0, cudaStream_t st,
                               const char* who) {
  STNLS_REQUIRE(njobs >= 1 && njobs <= LAYOUT_MAX_JOBS, "%s: 1..%d videos per launch", who, LAYOUT_MAX_JOBS);
  STNLS_REQUIRE(cl && planar && B && HD && t0 && Tloc && F && H && W && strides, "%s: null argument", who);
  LayoutJobs J;
  for (int j = 0; j < njobs; j++) {
    LayoutParams& P = J.job[j];
    P.cl = cl[j]; P.planar = planar[j]; P.HD = HD[j]; P.t0 = t0[j]; P.Tloc = Tloc[j]; P.F = F[j]; P.Fp = (F[j] + 3) / 4 * 4;
    P.H = H[j]; P.W = W[j]; P.sB = strides[4 * j]; P.sHD = strides[4 * j + 1]; P.sT = strides[4 * j + 2]; P.sC = strides[4 * j + 3];
    P.counts_h = counts_h ? counts_h[j] : nullptr; P.counts_w = counts_w ? counts_w[j] : nullptr;
    STNLS_REQUIRE((P.counts_h == nullptr) == (P.counts_w == nullptr), "%s: counts_h and counts_w go together", who);
    P.accumulate = accumulate ? accumulate[j] : 0;
    P.Tpitch = cl_tpitch ? cl_tpitch[j] : P.Tloc; P.tcl0 = cl_t0 ? cl_t0[j] : 0;
    STNLS_REQUIRE(P.tcl0 >= 0 && P.tcl0 + P.Tloc <= P.Tpitch, "%s: frames [%d, %d) outside the %d frames of the channels-last video",
                  who, P.tcl0, P.tcl0 + P.Tloc, P.Tpitch);
  }
  J.njobs = njobs;
  return launch_layout_jobs(pack, J, B, st, who);
}

extern "C" int stnls_pack_cl_multi(int njobs, float* const* dst_cl, const float* const* src, const int* B, const int* HD,
                                   const int* t0, const int* Tloc, const int* F, const int* H, const int* W,
                                   const int64_t* strides, const int* const* counts_h, const int* const* counts_w,
                                   const int* cl_tpitch, const int* cl_t0, stnls_stream_t stream) {
  return launch_layout_multi(true, njobs, dst_cl, const_cast<float* const*>(src), B, HD, t0, Tloc, F, H, W, strides,
                             counts_h, counts_w, nullptr, cl_tpitch, cl_t0, (cudaStream_t)stream, "pack_cl_multi");
}

extern "C" int stnls_unpack_cl_multi(int njobs, float* const* dst, const float* const* src_cl, const int* B, const int* HD,
                                     const int* t0, const int* Tloc, const int* F, const int* H, const int* W,
                                     const int64_t* strides, const int* accumulate, const int* cl_tpitch,
                                     const int* cl_t0, stnls_stream_t stream) {
  return launch_layout_multi(false, njobs, const_cast<float* const*>(src_cl), dst, B, HD, t0, Tloc, F, H, W, strides,
                             nullptr, nullptr, accumulate, cl_tpitch, cl_t0, (cudaStream_t)stream, "unpack_cl_multi");
}

extern "C" int stnls_add_inplace(float* dst, const float* src, int64_t n, stnls_stream_t stream) {
  STNLS_REQUIRE(dst && src, "add_inplace: null argument");
  if (n <= 0) return STNLS_OK;
  // the vector kernel needs 16-byte aligned operands (a frame slice of an odd-sized video is not): scalar otherwise
  const bool aligned = ((reinterpret_cast<uintptr_t>(dst) | reinterpret_cast<uintptr_t>(src)) & 15) == 0;
  int64_t nthr = aligned ? (n + 3) / 4 : n;
  STNLS_REQUIRE((nthr + 255) / 256 < 2147483647LL, "add_inplace: too many elements for one launch");
  unsigned blocks = (unsigned)((nthr + 255) / 256);
  if (aligned) add_inplace_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(dst, src, n);
  else add_inplace_scalar_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(dst, src, n);
  STNLS_CHECK_LAUNCH("add_inplace");
  return STNLS_OK;
}

static int seg_copy_impl(float* const* dst, const float* const* src, const int64_t* nfloat, const int* rows,
                         const int64_t* dst_pitch, const int64_t* src_pitch, int nseg, int accumulate, stnls_stream_t stream) {
  STNLS_REQUIRE(dst && src && nfloat && nseg >= 0 && nseg <= 8, "seg_copy: 0..8 segments");
  if (nseg == 0) return STNLS_OK;
  SegJobs J;
  int64_t most = 0;
  for (int i = 0; i < nseg; i++) {
    STNLS_REQUIRE(dst[i] && src[i] && nfloat[i] >= 0 && nfloat[i] % 4 == 0, "seg_copy: segment %d: null or not a multiple of 4 floats", i);
    STNLS_REQUIRE(((reinterpret_cast<uintptr_t>(dst[i]) | reinterpret_cast<uintptr_t>(src[i])) & 15) == 0,
                  "seg_copy: segment %d is not 16-byte aligned", i);
    J.dst[i] = dst[i]; J.src[i] = src[i]; J.n4[i] = nfloat[i] / 4;
    J.rows[i] = rows ? rows[i] : 1; J.dpitch[i] = rows ? dst_pitch[i] : 0; J.spitch[i] = rows ? src_pitch[i] : 0;
    STNLS_REQUIRE(J.rows[i] >= 0 && J.dpitch[i] % 4 == 0 && J.spitch[i] % 4 == 0, "seg_copy: segment %d: bad rows / pitch", i);
    if (J.rows[i] == 0) J.n4[i] = 0;
    most = J.n4[i] * J.rows[i] > most ? J.n4[i] * J.rows[i] : most;
  }
  J.nseg = nseg; J.accumulate = accumulate;
  if (most == 0) return STNLS_OK;
  int64_t bx = (most + 256 * 4 - 1) / (256 * 4);
  if (bx > 148 * 16) bx = 148 * 16;
  if (bx < 1) bx = 1;
  seg_copy_kernel<<<dim3((unsigned)bx, (unsigned)nseg), 256, 0, (cudaStream_t)stream>>>(J);
  STNLS_CHECK_LAUNCH("seg_copy");
  return STNLS_OK;
}

extern "C" int stnls_seg_copy(float* const* dst, const float* const* src, const int64_t* nfloat, int nseg, int accumulate,
                              stnls_stream_t stream) {
  return seg_copy_impl(dst, src, nfloat, nullptr, nullptr, nullptr, nseg, accumulate, stream);
}

extern "C" int stnls_seg_copy2d(float* const* dst, const float* const* src, const int64_t* nfloat, const int* rows,
                                const int64_t* dst_pitch, const int64_t* src_pitch, int nseg, int accumulate,
                                stnls_stream_t stream) {
  STNLS_REQUIRE(rows && dst_pitch && src_pitch, "seg_copy2d: null argument");
  return seg_copy_impl(dst, src, nfloat, rows, dst_pitch, src_pitch, nseg, accumulate, stream);
}
