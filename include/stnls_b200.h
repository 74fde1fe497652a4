/*
 * stnls_b200.h -- C ABI of the B200-native shifted non-local search + weighted patch sum.
 *
 * Drop-in boundary for the hot path of gauenk/stnls.  The reference binds
 * its kernels through a pybind11 module `stnls_cuda` whose free functions take torch::Tensor
 * (lib/csrc/pybind.cpp:30-57).  This header is what a maintainer would bind instead: plain
 * pointers, sizes and a cudaStream_t -- no torch types.  All pointers are DEVICE pointers, all
 * tensors fp32 unless stated; the caller owns every buffer (the reference's convention:
 * lib/stnls/search/impl/non_local_search.py:46-49,149-151).  Every call is asynchronous on
 * `stream`, never synchronises, and returns 0 on success or a negative STNLS_E* code; the last
 * error text is available from stnls_last_error().
 *
 * Internal video layout ("cl"): channels-last, one head per slab,
 *      (N = B*HD, Tloc, H, W, Fp)   Fp = F_head rounded up to a multiple of 4 (zero padded)
 * produced by stnls_pack_cl() from the reference's (B,HD,T,F,H,W) or (B,T,HD*F,H,W) tensors
 * (this replaces the copy `shape_vids` makes, lib/stnls/search/utils.py:178-190) and turned back
 * by stnls_unpack_cl().  Kernels read/write float4 channel vectors of that layout.
 *
 * Frame sharding (multi-GPU): a rank owns query frames [tq0, tq0+Tq) of a clip of T frames and
 * holds key/value frames [tk0, tk0+Tk) (its own plus the temporal halo).  Single GPU:
 * tq0=tk0=0, Tq=Tk=T.  All temporal index math uses the global T.
 */
#ifndef STNLS_B200_H
#define STNLS_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define STNLS_OK 0
#define STNLS_EINVAL (-1)   /* bad argument (shape / option outside the supported set) */
#define STNLS_ECUDA (-2)    /* CUDA launch error */

typedef void* stnls_stream_t; /* cudaStream_t */

/* ---- geometry shared by every entry point ------------------------------------------------ */
typedef struct {
  int B, HD;          /* batch, heads                                                        */
  int T;              /* global clip length                                                  */
  int tq0, Tq;        /* owned query frames [tq0, tq0+Tq)                                    */
  int tk0, Tk;        /* resident key/value frames [tk0, tk0+Tk)                             */
  int F, Fp;          /* channels per head, padded channels of the cl layout                 */
  int qH, qW, kH, kW; /* query / key frame sizes (equal for everything but strideQ searches) */
  int nH, nW;         /* query grid: (kH-1)/stride0+1, (kW-1)/stride0+1                      */
} stnls_geom;

/* ---- search options: mirrors the argument lists of
 *   non_local_search_int_forward / _bilin2d_forward  (lib/csrc/search/non_local_search.cpp:57-62,74-79)
 *   refinement_int_forward / _bilin2d_forward        (lib/csrc/search/refinement.cpp:39-44,56-62)
 * `ws,wt,K,Ks,wr` are explicit here (the reference infers them from tensor shapes,
 * non_local_search_int_kernel.cu:207-212). */
typedef struct {
  int ws, wt;          /* NLS: spatial / temporal window.  Refine: unused (0)                 */
  int Ks, wr;          /* Refine: number of input offsets and refinement window; NLS: 0       */
  int ps, K;           /* patch size; neighbours kept (K<=0: keep all L, unsorted)            */
  int stride0, strideQ, dilation;
  float stride1;
  int reflect_bounds, full_ws;
  int patch_offset;    /* 0 if use_adj else -(ps/2)  (impl/non_local_search.py:34)            */
  int off_Hq, off_Wq;
  int dist_type;       /* 0 = prod, 1 = l2  (lib/stnls/search/utils.py:225-227)               */
  int itype_int;       /* 1: int32 flows/inds (non-differentiable), 0: fp32                   */
  int anchor_self;     /* self_action in {"anchor","anchor_self"}                             */
  int HD_f;            /* heads of the flows tensor (hd % HD_f indexing)                      */
  int st_offset;       /* NLS: W_t - flows.size(3) in {0,1}                                   */
} stnls_search_opts;

const char* stnls_last_error(void);
int stnls_version(void);

/* ---- layout ------------------------------------------------------------------------------ */
/* planar (strided) -> cl.  src element (b,hd,t,c,h,w) at src[b*sB + hd*sHD + t*sT + c*sC + h*W + w];
 * frames t in [t0, t0+Tloc) are packed.  If counts_h/counts_w != NULL every pixel is divided by
 * (float)(counts_h[h]*counts_w[w]) on the way (gather_add.py:133-134 fused).                  */
int stnls_pack_cl(float* dst_cl, const float* src, int B, int HD, int t0, int Tloc, int F, int Fp,
                  int H, int W, int64_t sB, int64_t sHD, int64_t sT, int64_t sC,
                  const int* counts_h, const int* counts_w, stnls_stream_t stream);
/* cl -> planar (strided); accumulate != 0 adds into dst instead of overwriting                */
int stnls_unpack_cl(float* dst, const float* src_cl, int B, int HD, int t0, int Tloc, int F, int Fp,
                    int H, int W, int64_t sB, int64_t sHD, int64_t sT, int64_t sC,
                    int accumulate, stnls_stream_t stream);

/* several videos in ONE launch (q, k, v of a step; the three gradients on the way back): the per-video arguments of
 * stnls_pack_cl / stnls_unpack_cl as HOST arrays of njobs <= 4 entries; strides holds (sB, sHD, sT, sC) per video;
 * Fp = F rounded up to 4; counts_h / counts_w / accumulate may be NULL (no division / overwrite).  Same copy as
 * shape_vids (lib/stnls/search/utils.py:178-190) for all inputs of a search + aggregation step at once.              */
int stnls_pack_cl_multi(int njobs, float* const* dst_cl, const float* const* src, const int* B, const int* HD,
                        const int* t0, const int* Tloc, const int* F, const int* H, const int* W,
                        const int64_t* strides, const int* const* counts_h, const int* const* counts_w,
                        const int* cl_tpitch, const int* cl_t0, stnls_stream_t stream);
int stnls_unpack_cl_multi(int njobs, float* const* dst, const float* const* src_cl, const int* B, const int* HD,
                          const int* t0, const int* Tloc, const int* F, const int* H, const int* W,
                          const int64_t* strides, const int* accumulate, const int* cl_tpitch, const int* cl_t0,
                          stnls_stream_t stream);
/* cl_tpitch / cl_t0 (NULL: Tloc / 0): the channels-last video of job j holds cl_tpitch[j] frames per (b, head) and the
 * Tloc[j] frames moved start at its frame cl_t0[j] -- a rank packs its own frames straight into the owned + halo
 * buffer its peers read, and unpacks the owned slice of an owned + halo gradient.                                  */

/* ---- search forward: kernel + anchor_self + top-K fused ------------------------------------
 * replaces  stnls_cuda.non_local_search_{int,bilin2d}_forward + stnls_cuda.anchor_self + stnls.nn.topk
 *           (impl/non_local_search.py:57-126) and
 *           stnls_cuda.refinement_{int,bilin2d}_forward + anchor_self_refine + topk (impl/refinement.py:56-98)
 *  q_cl (N,Tq,qH,qW,Fp)  k_cl (N,Tk,kH,kW,Fp)
 *  flows: NLS    (B,HD_f,Tq,W_t-st_offset,2,nH,nW)   int32 or fp32
 *         Refine (B,HD_f,Tq,nH,nW,Ks,3)
 *  dists (B,HD,Tq,nH,nW,Kout)   inds (B,HD,Tq,nH,nW,Kout,3) int32/fp32
 *      Kout = K>0 ? min(K,L) : L ,  L = (2wt+1)*ws*ws  or  Ks*wr*wr ;  fused top-K needs K<=32
 *  kselect (B,HD,Tq,nH,nW,Kout) int32 or NULL ; reflect (B,HD_f,Tq,nH,nW,Ks,2) uint8, pre-zeroed, or NULL
 */
int stnls_search_forward(const float* q_cl, const float* k_cl, const void* flows,
                         float* dists, void* inds, int* kselect, uint8_t* reflect,
                         const stnls_geom* g, const stnls_search_opts* o, stnls_stream_t stream);

/* ---- search backward -------------------------------------------------------------------------
 * replaces  non_local_search_int_vid_backward / _bilin2d_vid_backward / _bilin2d_vidflows_backward
 *           (lib/csrc/search/non_local_search.cpp:91-97,112-118,132-138) and
 *           refinement_bilin2d_vidflows_backward (lib/csrc/search/refinement.cpp:77-86)
 *  gq_cl (N,Tq,qH,qW,Fp), gk_cl (N,Tk,kH,kW,Fp): pre-zeroed accumulators
 *  grad_flows: NULL, or pre-zeroed, same shape as flows (NLS layout or refine layout)
 *  grad_dists (B,HD,Tq,nH,nW,K); grad_inds (…,K,3) (only read when grad_flows != NULL)
 */
int stnls_search_backward(float* gq_cl, float* gk_cl, float* grad_flows,
                          const float* q_cl, const float* k_cl, const float* flows,
                          const float* grad_dists, const float* grad_inds, const void* inds,
                          const int* kselect, const uint8_t* reflect, int K,
                          const stnls_geom* g, const stnls_search_opts* o, stnls_stream_t stream);

/* search backward with the normalise step's backward folded in (SoftmaxNormalize, lib/stnls/normz/softmax.py:21-42,
 * between the search and the aggregation): instead of grad_dists the kernel takes the gradient of the weights and the
 * weights (B,HD,Tq,nH,nW,K) and uses grad_dists[k] = scale * w[k] * (gw[k] - sum_j w[j] gw[j]) -- the arithmetic of
 * stnls_softmax_backward, bit for bit.  Everything else as stnls_search_backward.                                  */
int stnls_search_backward_sm(float* gq_cl, float* gk_cl, float* grad_flows,
                             const float* q_cl, const float* k_cl, const float* flows,
                             const float* grad_weights, const float* weights, float scale,
                             const float* grad_inds, const void* inds,
                             const int* kselect, const uint8_t* reflect, int K,
                             const stnls_geom* g, const stnls_search_opts* o, stnls_stream_t stream);

/* ---- weighted patch sum (== NonLocalGatherAdd) ------------------------------------------------
 * replaces  gather_add_forward / gather_add_int_backward / gather_add_bilin2d_backward
 *           (lib/csrc/agg/gather_add.cpp:53-57,68-73,86-92)
 *  v_cl (N,Tk,H,W,Fp); weights (B,HD,Tq,nH,nW,K); flows (B,HD,Tq,nH,nW,K,3) int32/fp32
 *  out_cl (N,Tq,H,W,Fp): written (gather form, already divided by the patch-overlap counts)
 *  counts_h (H), counts_w (W): separable overlap counts, counts[h][w] = counts_h[h]*counts_w[w]
 */
typedef struct {
  int ps, K, stride0, dilation, reflect_bounds, patch_offset, itype_int;
} stnls_agg_opts;

int stnls_wpsum_forward(float* out_cl, const float* v_cl, const float* weights, const void* flows,
                        const int* counts_h, const int* counts_w,
                        const stnls_geom* g, const stnls_agg_opts* o, stnls_stream_t stream);
/* the same with the normalise step folded in: weights = softmax(scale * dists[..., :K]) is computed
 * inside the aggregation kernel -- the arithmetic of stnls_softmax_forward, bit for bit -- and WRITTEN to `weights`
 * (B,HD,Tq,nH,nW,K) for the backward pass; rows of dists hold Kin >= K entries.  *launches (may be NULL) = kernels
 * run: 1, or 2 where the fused kernel does not apply (ps != 3, dilation != 1, more than 32 channels per head) and
 * the normalise kernel runs first.                                                                               */
int stnls_wpsum_forward_sm(float* out_cl, float* weights, const float* v_cl, const float* dists, int Kin, float scale,
                           const void* flows, const int* counts_h, const int* counts_w,
                           const stnls_geom* g, const stnls_agg_opts* o, stnls_stream_t stream, int* launches);
/*  gout_cl (N,Tq,H,W,Fp): incoming gradient already divided by counts (stnls_pack_cl does it)
 *  gv_cl (N,Tk,H,W,Fp) pre-zeroed; grad_weights (B,HD,Tq,nH,nW,K) pre-zeroed;
 *  grad_flows (…,K,3) pre-zeroed or NULL (int)                                                  */
int stnls_wpsum_backward(float* gv_cl, float* grad_weights, float* grad_flows,
                         const float* gout_cl, const float* v_cl, const float* weights,
                         const void* flows, const stnls_geom* g, const stnls_agg_opts* o,
                         stnls_stream_t stream);

/* ---- NonLocalGather (K-stacked aggregation) ---------------------------------------
 * replaces  gather_int_forward / gather_bilin2d_forward / gather_int_backward / gather_bilin2d_backward
 *           (lib/csrc/agg/gather.cpp; kernels lib/csrc/agg/gather_{int,bilin2d}_kernel.cu, driven by
 *           lib/stnls/agg/gather.py:70-276).  Same arguments as the weighted patch sum, but the sum over the
 *           K neighbours is NOT taken: stack_cl / gstack_cl are (N*K, Tq, H, W, Fp), slab n*K + k holding
 *           neighbour k -- the cl form of the reference's (B,HD,K,T,F,H,W) stack.                        */
int stnls_gather_forward(float* stack_cl, const float* v_cl, const float* weights, const void* flows,
                         const int* counts_h, const int* counts_w,
                         const stnls_geom* g, const stnls_agg_opts* o, stnls_stream_t stream);
int stnls_gather_backward(float* gv_cl, float* grad_weights, float* grad_flows,
                          const float* gstack_cl, const float* v_cl, const float* weights,
                          const void* flows, const stnls_geom* g, const stnls_agg_opts* o,
                          stnls_stream_t stream);

/* ---- search_flow (accumulated per-query flows) ---------------------------------------
 * replaces  search_flow_forward / search_flow_backward (lib/csrc/nn/search_flow.cpp:29-58, kernels
 *           lib/csrc/nn/search_flow_kernel.cu), driven by lib/stnls/nn/search_flow.py:24-96.
 *  fflow, bflow (B,T,2,H,W): optical flow t->t+1 / t->t-1, channel 0 = dw, 1 = dh
 *  flows (B,T,W_t-1,2,nH,nW), W_t = min(2*wt+1,T): written (forward) / read (backward)
 *  grad_fflow, grad_bflow (B,T,2,H,W): pre-zeroed accumulators                                       */
int stnls_search_flow_forward(float* flows, const float* fflow, const float* bflow, int B, int T,
                              int H, int W, int wt, int stride0, stnls_stream_t stream);
int stnls_search_flow_backward(float* grad_fflow, float* grad_bflow, const float* grad_flows,
                               const float* fflow, const float* bflow, const float* flows, int B,
                               int T, int H, int W, int wt, int stride0, stnls_stream_t stream);

/* ---- fused normalise step --------------------------------------------------------
 * replaces  SoftmaxNormalize (lib/stnls/normz/softmax.py:21-42) / `softmax(10*dists,-1)` (README.md:51):
 *   weights[r, k] = softmax_k(scale * dists[r, k]),  k < K;  dists rows have Kin >= K entries (k_agg slice),
 *   `scale` carries the sign (negative for l2 distances).  K <= 64.
 *  backward: grad_dists (rows, Kin) fully written (zeros for k >= K).                                    */
int stnls_softmax_forward(float* weights, const float* dists, int64_t rows, int K, int Kin,
                          float scale, stnls_stream_t stream);
int stnls_softmax_backward(float* grad_dists, const float* grad_weights, const float* weights,
                           int64_t rows, int K, int Kin, float scale, stnls_stream_t stream);

/* ---- per-frame anchoring of the un-reduced candidate list -----------------------------------------
 * replaces  stnls.nn.anchor_self_time (lib/stnls/nn/anchor_self.py:69-79 -> anchor_self_time_forward_cuda,
 * lib/csrc/nn/anchor_self_kernel.cu:187-360), the step behind self_action = "anchor_each" and
 * "anchor_and_remove_ref_frame" (lib/stnls/search/impl/non_local_search.py:96-110).  In place:
 *  dists (B,HD,Q,G,S) f32, inds (B,HD,Q,G,S,3) i32|f32 (keep-all output of stnls_search_forward, Q = T*nH*nW,
 *  G = 2wt+1 or 2wt frame groups, S = ws*ws), flows (B,HD_f,T,nflow,2,nH,nW) of the index type.           */
int stnls_anchor_self_time(float* dists, void* inds, const void* flows, int B, int HD, int HD_f, int T,
                           int G, int S, int nflow, int wt, int stride0, int qH, int qW, int kH, int kW,
                           int itype_int, stnls_stream_t stream);

/* separable patch-overlap counts (host helper, no GPU): counts_h[h] = #{(nh,pi): nh*stride0 +
 * dilation*(pi+patch_offset) == h}; equals what gather_add_forward_kernel counts atomically
 * (lib/csrc/agg/gather_add_kernel.cu:94-97).  Returns the minimum count (0 => reference asserts). */
int stnls_overlap_counts(int* counts, int H, int nH, int ps, int stride0, int dilation, int patch_offset);

/* ---- halo add: dst[i] += src[i]  (multi-GPU gradient halo reduce); any alignment ------------ */
int stnls_add_inplace(float* dst, const float* src, int64_t n, stnls_stream_t stream);

/* ---- halo staging / pull: up to 8 contiguous segments copied (accumulate == 0) or added with vector REDs
 * (accumulate != 0) in one launch; dst[i], src[i] 16-byte aligned, nfloat[i] a multiple of 4.  src may be memory of a
 * peer GPU mapped into this process (NVLink peer loads).  The arrays are HOST arrays of nseg entries.          */
int stnls_seg_copy(float* const* dst, const float* const* src, const int64_t* nfloat, int nseg, int accumulate,
                   stnls_stream_t stream);

/* the same with rows: segment i is rows[i] runs of nfloat[i] floats, dst_pitch[i] / src_pitch[i] floats apart (multiples
 * of 4) -- the frames [lo, hi) of every (batch, head) slab of a channels-last video whose slabs have different lengths
 * on the two sides (a peer's owned + halo buffer and this rank's).                                                */
int stnls_seg_copy2d(float* const* dst, const float* const* src, const int64_t* nfloat, const int* rows,
                     const int64_t* dst_pitch, const int64_t* src_pitch, int nseg, int accumulate, stnls_stream_t stream);

/* ---- measurement aid: fp32 FMA-pipe microbenchmark (no reference counterpart) -------------------
 * `blocks` CTAs of 256 threads, each thread 16 independent FFMA chains of `iters` steps; flops per launch =
 * 2*16*iters*256*blocks.  bench.py times it with CUDA events for the fp32 roofline denominator. */
int stnls_ffma_peak(float* sink, int blocks, int iters, stnls_stream_t stream);

/* ---- allocation-free forward search on planar videos ---------------------------------------------
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same result as  stnls_pack_cl(q) + stnls_pack_cl(k) + stnls_search_forward(..., kselect = NULL, reflect = NULL),
 * but the two channels-last copies live in a device array built into the library (STNLS_NOALLOC_MAX_BYTES, held by
 * every GPU that loads the library until exit) instead of caller-allocated buffers.  No cudaMalloc / cudaFree /
 * allocating routine is called.
 *  q (B, T, HD*F, qH, qW), k (B, T, HD*F, kH, kW): contiguous planar videos of the whole clip; frames
 *  [tq0, tq0+Tq) of q and [tk0, tk0+Tk) of k are used.  flows / dists / inds: as for stnls_search_forward.
 *  Limit: round256(B*HD*Tq*qH*qW*Fp*4) + round256(B*HD*Tk*kH*kW*Fp*4) <= STNLS_NOALLOC_MAX_BYTES, else STNLS_EINVAL.
 *  The array is shared by every call on a GPU and keeps the last call's contents (each call overwrites all it reads,
 *  nothing needs clearing): calls that overlap on one GPU -- two streams, two threads -- RACE; serialise them.
 *  The device that owns the buffers must be current, as for every other entry point.                              */
#define STNLS_NOALLOC_MAX_BYTES (32u << 20)
int stnls_search_forward_planar_noalloc(const float* q, const float* k, const void* flows, float* dists, void* inds,
                                        const stnls_geom* g, const stnls_search_opts* o, stnls_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif
