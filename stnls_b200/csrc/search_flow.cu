// search_flow: optical flows between consecutive frames -> accumulated per-query offsets for every slot
// of the temporal search window (the step immediately before the search).
// Replaces
//   search_flow_forward_kernel   lib/csrc/nn/search_flow_kernel.cu:17-88  (+ shared_flows.cu:39-72)
//   search_flow_backward_kernel  lib/csrc/nn/search_flow_kernel.cu:131-250 (+ shared_flows.cu:108-290)
// fflow/bflow (B,T,2,H,W) (channel 0 = dw, 1 = dh); flows (B,T,W_t-1,2,nH,nW), W_t = min(2wt+1,T).
// The tensors are two-channel and small next to the videos, so the kernels are latency-bound chains of
// dependent bilinear reads: one thread per (batch, query) forward, one per (batch, query, slot) backward,
// indexed so that a warp walks neighbouring pixels of one row (coalesced reads of the flow planes).
#include "common.cuh"

namespace {

struct FlowParams {
  float* flows;        // fwd: out ; bwd: in
  const float* fflow;
  const float* bflow;
  const float* gflows;
  float* gfflow;
  float* gbflow;
  int B, T, H, W, nH, nW, wt, stride0, W_t;
  int64_t nthreads;
};

__device__ __forceinline__ int clampi(int v, int lim) { return min(max(v, 0), lim - 1); }

// one hop: centre += bilinear(flow)(centre)  (shared_flows.cu:39-72: floor taps, weights from the un-reflected
// tap, taps reflected once).  A tap still outside the frame after one reflection is an out-of-bounds read in
// the reference; it is clamped here.
__device__ __forceinline__ void hop(float& hc, float& wc, const float* flow, int H, int W) {
  const float h0 = hc, w0 = wc;
  const int64_t plane = (int64_t)H * W;
#pragma unroll
  for (int i = 0; i < 2; i++) {
#pragma unroll
    for (int j = 0; j < 2; j++) {
      int hj = __float2int_rd(h0 + i), wj = __float2int_rd(w0 + j);
      const float weight = fmaxf(0.f, 1 - fabsf(hj - h0)) * fmaxf(0.f, 1 - fabsf(wj - w0));
      hj = clampi(bnd(hj, H), H);
      wj = clampi(bnd(wj, W), W);
      wc = fmaf(weight, __ldg(flow + (int64_t)hj * W + wj), wc);
      hc = fmaf(weight, __ldg(flow + plane + (int64_t)hj * W + wj), hc);
    }
  }
}

__global__ void __launch_bounds__(256) search_flow_fwd_kernel(const __grid_constant__ FlowParams P) {
  int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= P.nthreads) return;
  const int nw = (int)(idx % P.nW); idx /= P.nW;
  const int nh = (int)(idx % P.nH); idx /= P.nH;
  const int ti = (int)(idx % P.T);
  const int b = (int)(idx / P.T);
  const int H = P.H, W = P.W;
  const float h_ref = (float)((nh * P.stride0) % H), w_ref = (float)((nw * P.stride0) % W);
  const int t_max = time_tmax(ti, P.T, P.wt);
  const int64_t fs = 2LL * H * W;                     // one flow frame
  const int64_t qplane = (int64_t)P.nH * P.nW;
  float* out = P.flows + ((int64_t)b * P.T + ti) * (P.W_t - 1) * 2 * qplane + (int64_t)nh * P.nW + nw;
  float hc = h_ref, wc = w_ref;
  for (int si = 1; si < P.W_t; si++) {
    int tj = ti + si;
    const bool swap = (ti + si - 1) == t_max;       // first backward slot: restart from the query pixel
    tj = (tj > t_max) ? t_max - si : tj;
    if (swap) { hc = h_ref; wc = w_ref; }
    const float* flow = (tj > ti) ? P.fflow + ((int64_t)b * P.T + (tj - 1)) * fs
                                  : P.bflow + ((int64_t)b * P.T + (tj + 1)) * fs;
    hop(hc, wc, flow, H, W);
    out[((int64_t)(si - 1) * 2 + 1) * qplane] = hc - h_ref;
    out[((int64_t)(si - 1) * 2 + 0) * qplane] = wc - w_ref;
  }
}

// Jacobian update of one later hop (shared_flows.cu:151-290): dA/dF <- dA/dF + dFlow(prop) * S * dA/dF
__device__ __forceinline__ void update_weights(float* dAdF0, float* dAdF1, float p1, float p2, int H, int W,
                                               const float* flow) {
  float dF00 = 0.f, dF01 = 0.f, dF10 = 0.f, dF11 = 0.f;
  const int sH = inb(p1, H) ? 1 : -1, sW = inb(p2, W) ? 1 : -1;
  p1 = bnd(p1, H);
  p2 = bnd(p2, W);
  const int64_t plane = (int64_t)H * W;
#pragma unroll
  for (int ix = 0; ix < 2; ix++) {
#pragma unroll
    for (int jx = 0; jx < 2; jx++) {
      const int i0 = __float2int_rz(ix == 0 ? floorf(p1) : floorf(p1) + 1);
      const float gH = fmaxf(0.f, 1 - fabsf(i0 - p1));
      const int i1 = __float2int_rz(jx == 0 ? floorf(p2) : floorf(p2) + 1);
      const float gW = fmaxf(0.f, 1 - fabsf(i1 - p2));
      if (!inb(i0, H) || !inb(i1, W)) continue;
      const float vW = __ldg(flow + (int64_t)i0 * W + i1), vH = __ldg(flow + plane + (int64_t)i0 * W + i1);
      dF00 += (jx == 0) ? -gH * vW : gH * vW;
      dF01 += (ix == 0) ? -gW * vW : gW * vW;
      dF10 += (jx == 0) ? -gH * vH : gH * vH;
      dF11 += (ix == 0) ? -gW * vH : gW * vH;
    }
  }
  const float a0 = dAdF0[0], a1 = dAdF0[1], b0 = dAdF1[0], b1 = dAdF1[1];
  dAdF0[0] += dF00 * sW * a0 + dF01 * sH * a1;
  dAdF0[1] += dF10 * sW * a0 + dF11 * sH * a1;
  dAdF1[0] += dF00 * sW * b0 + dF01 * sH * b1;
  dAdF1[1] += dF10 * sW * b0 + dF11 * sH * b1;
}

// scatter to the 4 taps of the first hop (shared_flows.cu:108-141)
__device__ __forceinline__ void assign_bilin2d(const float* dAdF0, const float* dAdF1, float gv0, float gv1,
                                               float r1, float r2, int H, int W, float* g_flow) {
  const int64_t plane = (int64_t)H * W;
  const float v0 = dAdF0[0] * gv0 + dAdF0[1] * gv1, v1 = dAdF1[0] * gv0 + dAdF1[1] * gv1;
#pragma unroll
  for (int ix = 0; ix < 2; ix++) {
#pragma unroll
    for (int jx = 0; jx < 2; jx++) {
      int i0 = __float2int_rd(r1 + ix);
      const float gH = fmaxf(0.f, 1 - fabsf(i0 - r1));
      int i1 = __float2int_rd(r2 + jx);
      const float gW = fmaxf(0.f, 1 - fabsf(i1 - r2));
      const float w = gH * gW;
      i0 = clampi(bnd(i0, H), H);
      i1 = clampi(bnd(i1, W), W);
      atomicAdd(g_flow + (int64_t)i0 * W + i1, w * v0);
      atomicAdd(g_flow + plane + (int64_t)i0 * W + i1, w * v1);
    }
  }
}

__global__ void __launch_bounds__(256) search_flow_bwd_kernel(const __grid_constant__ FlowParams P) {
  int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= P.nthreads) return;
  const int nw = (int)(idx % P.nW); idx /= P.nW;
  const int nh = (int)(idx % P.nH); idx /= P.nH;
  const int si = (int)(idx % (P.W_t - 1)) + 1; idx /= (P.W_t - 1);   // slot W_t of the reference's grid is a no-op
  const int ti = (int)(idx % P.T);
  const int b = (int)(idx / P.T);
  const int H = P.H, W = P.W, T = P.T;
  const int rh = (nh * P.stride0) % H, rw = (nw * P.stride0) % W;
  const int t_shift = min(0, ti - P.wt) + max(0, ti + P.wt - (T - 1));
  const int t_min = max(ti - P.wt - t_shift, 0), t_max = min(T - 1, ti + P.wt - t_shift);
  int tj = ti + si;
  tj = (tj > t_max) ? t_max - si : tj;
  const bool isFwd = tj > ti;
  const int t_flow = isFwd ? tj - 1 : tj + 1;
  const int64_t fs = 2LL * H * W;
  const float* flow = (isFwd ? P.fflow : P.bflow) + (int64_t)b * T * fs;
  float* g_flow = (isFwd ? P.gfflow : P.gbflow) + (int64_t)b * T * fs;
  const int t_inc = isFwd ? 1 : -1;
  const int t_end = isFwd ? (t_max - tj + 1) : (tj - t_min + 1);
  const bool adjacent = abs(ti - tj) == 1;
  const int64_t qplane = (int64_t)P.nH * P.nW;
  const int64_t qo = (int64_t)nh * P.nW + nw;
  const float* fl = P.flows + ((int64_t)b * T + ti) * (P.W_t - 1) * 2 * qplane + qo;
  const float* gf = P.gflows + ((int64_t)b * T + ti) * (P.W_t - 1) * 2 * qplane + qo;
  float r1 = (float)rh, r2 = (float)rw;
  if (!adjacent) {
    r1 = rh + fl[((int64_t)(si - 2) * 2 + 1) * qplane];
    r2 = rw + fl[((int64_t)(si - 2) * 2 + 0) * qplane];
  }
  float dAdF0[2], dAdF1[2];
  for (int tx = 0; tx < t_end; tx++) {
    const float gv_W = gf[((int64_t)(si - 1 + tx) * 2 + 0) * qplane];
    const float gv_H = gf[((int64_t)(si - 1 + tx) * 2 + 1) * qplane];
    if (tx == 0) {
      dAdF0[0] = 1; dAdF0[1] = 0; dAdF1[0] = 0; dAdF1[1] = 1;
    } else {
      const float p1 = (float)rh + fl[((int64_t)(si + tx - 2) * 2 + 1) * qplane];
      const float p2 = (float)rw + fl[((int64_t)(si + tx - 2) * 2 + 0) * qplane];
      update_weights(dAdF0, dAdF1, p1, p2, H, W, flow + (int64_t)(t_flow + t_inc * tx) * fs);
    }
    assign_bilin2d(dAdF0, dAdF1, gv_W, gv_H, r1, r2, H, W, g_flow + (int64_t)t_flow * fs);
  }
}

int flow_check(const char* who, int B, int T, int H, int W, int wt, int stride0) {
  STNLS_REQUIRE(B > 0 && T > 0 && H > 0 && W > 0, "%s: empty tensor", who);
  STNLS_REQUIRE(wt >= 0 && stride0 > 0, "%s: need wt >= 0 and stride0 > 0", who);
  return STNLS_OK;
}

FlowParams flow_params(int B, int T, int H, int W, int wt, int stride0) {
  FlowParams P = {};
  P.B = B; P.T = T; P.H = H; P.W = W; P.wt = wt; P.stride0 = stride0;
  P.nH = (H - 1) / stride0 + 1; P.nW = (W - 1) / stride0 + 1;
  P.W_t = (2 * wt + 1 < T) ? 2 * wt + 1 : T;
  return P;
}

}  // namespace

extern "C" int stnls_search_flow_forward(float* flows, const float* fflow, const float* bflow, int B, int T,
                                         int H, int W, int wt, int stride0, stnls_stream_t stream) {
  if (int rc = flow_check("search_flow_forward", B, T, H, W, wt, stride0)) return rc;
  STNLS_REQUIRE(flows && fflow && bflow, "search_flow_forward: null argument");
  FlowParams P = flow_params(B, T, H, W, wt, stride0);
  if (P.W_t <= 1) return STNLS_OK;
  P.flows = flows; P.fflow = fflow; P.bflow = bflow;
  P.nthreads = (int64_t)B * T * P.nH * P.nW;
  STNLS_REQUIRE(P.nthreads / 256 < 2147483647LL, "search_flow_forward: too many queries for one launch");
  search_flow_fwd_kernel<<<(unsigned)((P.nthreads + 255) / 256), 256, 0, (cudaStream_t)stream>>>(P);
  STNLS_CHECK_LAUNCH("search_flow_forward");
  return STNLS_OK;
}

extern "C" int stnls_search_flow_backward(float* grad_fflow, float* grad_bflow, const float* grad_flows,
                                          const float* fflow, const float* bflow, const float* flows, int B,
                                          int T, int H, int W, int wt, int stride0, stnls_stream_t stream) {
  if (int rc = flow_check("search_flow_backward", B, T, H, W, wt, stride0)) return rc;
  STNLS_REQUIRE(grad_fflow && grad_bflow && grad_flows && fflow && bflow && flows,
                "search_flow_backward: null argument");
  FlowParams P = flow_params(B, T, H, W, wt, stride0);
  if (P.W_t <= 1) return STNLS_OK;
  P.flows = const_cast<float*>(flows); P.fflow = fflow; P.bflow = bflow;
  P.gflows = grad_flows; P.gfflow = grad_fflow; P.gbflow = grad_bflow;
  P.nthreads = (int64_t)B * T * (P.W_t - 1) * P.nH * P.nW;
  STNLS_REQUIRE(P.nthreads / 256 < 2147483647LL, "search_flow_backward: too many threads for one launch");
  search_flow_bwd_kernel<<<(unsigned)((P.nthreads + 255) / 256), 256, 0, (cudaStream_t)stream>>>(P);
  STNLS_CHECK_LAUNCH("search_flow_backward");
  return STNLS_OK;
}
