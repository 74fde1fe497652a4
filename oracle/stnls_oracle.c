/*
 * oracle/stnls_oracle.c  --  TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement (plain C + OpenMP) of the reference's shifted non-local search,
 * refinement search and weighted-patch-sum (gather_add) kernels.  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it.
 * The product path (stnls_b200/) never imports or links anything under oracle/.
 *
 * Every function cites the reference file:line (under /root/reference/) it restates.
 * Tensor layouts are the reference's: vid (B,HD,T,F,H,W), flows (B,HD_f,T,W_t',2,nH,nW),
 * full candidate tensors dists (B,HD,Q,W_t,ws,ws), inds (...,3).
 *
 * Floating-point notes: nvcc contracts a*b+c into FMA by default, so the places where the
 * reference kernels accumulate (`dist += pix0*pix1`, `pix += w*vid`) use fmaf() here; gcc is run
 * with -ffp-contract=off so nothing else is contracted.
 *
 * Parity status: pinned on the GPU box against the reference's own kernels (oracle/_ref, see
 * oracle/build_ref.py and tests/test_ref_parity.py) and against the literal known-answer vectors
 * of dev/trange.py:33-41 and dev/ws_offset.py:37-41 (tests/test_oracle_kat.py).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define IDX6(a,b,c,d,e,f,B_,C_,D_,E_,F_) ((((((int64_t)(a)*(B_)+(b))*(C_)+(c))*(D_)+(d))*(E_)+(e))*(F_)+(f))

/* lib/csrc/shared_kernel.cu:49-60  bounds<int> (reflect once) */
static inline int bounds_i(int v, int lim){
  if (v < 0) return -v;
  if (v > lim-1) return 2*(lim-1)-v;
  return v;
}
/* lib/csrc/shared_kernel.cu:49-60  bounds<float> */
static inline float bounds_f(float v, int lim){
  if (v < 0) return -v;
  if (v > (float)(lim-1)) return (float)(2*(lim-1)) - v;
  return v;
}
/* lib/csrc/shared_kernel.cu:143-153 check_interval / check_bound */
static inline int inb_i(int v, int lim){ return (v >= 0) && (v <= lim-1); }
static inline int inb_f(float v, int lim){ return (v >= 0) && (v <= (float)(lim-1)); }

/* lib/csrc/shared_kernel.cu:76-98 get_pixel_loc<int> */
static inline void pixel_loc(int* pix, int qi, int stride0, int nW, int nHW, int H, int W){
  int tmp = qi;
  pix[0] = tmp / nHW;
  tmp = tmp - pix[0]*nHW;
  int nh = tmp / nW;
  pix[1] = (nh*stride0) % H;
  tmp = tmp - nh*nW;
  pix[2] = ((tmp % nW) * stride0) % W;
}

/* lib/csrc/shared_kernel.cu:307-312 set_time_range */
static inline int time_tmax(int ti, int T, int wt){
  int a = ti - wt; if (a > 0) a = 0;
  int b = ti + wt - (T-1); if (b < 0) b = 0;
  int t_shift = a + b;
  int t_max = ti + wt - t_shift; if (t_max > T-1) t_max = T-1;
  return t_max;
}
/* exported for the known-answer test of dev/trange.py:33-41 */
void oracle_time_window(int ti, int wt, int T, int* out){
  int W_t = 2*wt+1;
  int t_max = time_tmax(ti,T,wt);
  for (int si=0; si<W_t; si++){
    int tj = ti + si;                      /* non_local_search_int_kernel.cu:123-124 */
    tj = (tj > t_max) ? t_max - si : tj;
    out[si] = tj;
  }
}

/* lib/csrc/shared_kernel.cu:248-287 set_search_offsets<int> */
static inline int search_off_i(int a, int s1, int wsHalf, int ws, int lim, int full_ws){
  if (!full_ws) return wsHalf;
  int off = wsHalf;
  if ((a - s1*wsHalf) < 0) off = (int)floor(a/(1.*s1));
  int aMax = a + s1*((ws-1) - off);
  if (aMax > (lim-1)) off = (int)ceil((a - (lim-1))/(1.*s1) + (ws-1));
  return off;
}
/* lib/csrc/shared_kernel.cu:248-287 set_search_offsets<float> */
static inline float search_off_f(float a, float s1, float wsHalf, int ws, int lim, int full_ws){
  if (!full_ws) return wsHalf;
  float off = wsHalf;
  if (fmaf(-s1,wsHalf,a) < 0) off = (float)floor(a/(1.*s1));
  float aMax = fmaf(s1,((float)(ws-1) - off),a);
  if (aMax > (float)(lim-1)) off = (float)ceil((a - (float)(lim-1))/(1.*s1) + (ws-1));
  return roundf(off);
}
/* exported for the known-answer test of dev/ws_offset.py:37-41 */
void oracle_search_offset_f(float hi, float wi, float s1, int ws, int H, int W, float* out){
  float wsHalf = (float)((ws-1)/2);
  out[0] = search_off_f(hi,s1,wsHalf,ws,H,1);
  out[1] = search_off_f(wi,s1,wsHalf,ws,W,1);
}
void oracle_search_offset_i(int hi, int wi, int s1, int ws, int H, int W, int* out){
  int wsHalf = (ws-1)/2;
  out[0] = search_off_i(hi,s1,wsHalf,ws,H,1);
  out[1] = search_off_i(wi,s1,wsHalf,ws,W,1);
}

/* lib/csrc/shared_kernel.cu:738-771 bilin2d_interpolate; vid is one (H,W) plane */
static inline float bilin2d_interp(const float* vid, float hi, float wi, int H, int W){
  float pix = 0;
  for (int ix=0; ix<2; ix++){
    for (int jx=0; jx<2; jx++){
      int h_i = (int)truncf(hi+ix);
      float w = fmaxf(0.f, 1 - fabsf(h_i-hi));
      int w_i = (int)truncf(wi+jx);
      w = w*fmaxf(0.f, 1 - fabsf(w_i-wi));
      if (!inb_i(h_i,H)) continue;
      if (!inb_i(w_i,W)) continue;
      pix = fmaf(w, vid[(int64_t)h_i*W+w_i], pix);
    }
  }
  return pix;
}
/* lib/csrc/shared_kernel.cu:829-860 bilin2d_assign (atomic scatter) */
static inline void bilin2d_assign(float val, float hi, float wi, int H, int W, float* tensor){
  for (int ix=0; ix<2; ix++){
    for (int jx=0; jx<2; jx++){
      int h_i = (int)truncf(hi+ix);
      float w = fmaxf(0.f, 1 - fabsf(h_i-hi));
      int w_i = (int)truncf(wi+jx);
      w = w*fmaxf(0.f, 1 - fabsf(w_i-wi));
      if (!inb_i(h_i,H)) continue;
      if (!inb_i(w_i,W)) continue;
      float add = w*val;
      #pragma omp atomic
      tensor[(int64_t)h_i*W+w_i] += add;
    }
  }
}
/* lib/csrc/shared_kernel.cu:916-964 update_dFlows */
static inline void update_dflows(float* dF, float dD, float hi, float wi, int H, int W,
                                 int signH, int signW, const float* vid1){
  for (int ix=0; ix<2; ix++){
    for (int jx=0; jx<2; jx++){
      int h_i = (int)truncf(hi+ix);
      float gH = fmaxf(0.f, 1 - fabsf(h_i-hi));
      int w_i = (int)truncf(wi+jx);
      float gW = fmaxf(0.f, 1 - fabsf(w_i-wi));
      gH = (jx==0) ? -gH : gH;
      gW = (ix==0) ? -gW : gW;
      if (!inb_i(h_i,H)) continue;
      if (!inb_i(w_i,W)) continue;
      float pix = vid1[(int64_t)h_i*W+w_i];
      int kx = ix*4 + jx*2;
      dF[kx]   += gH*pix*dD*signW;
      dF[kx+1] += gW*pix*dD*signH;
    }
  }
}

typedef struct {
  int B, HD, HD_f, T, F, qH, qW, kH, kW, nH, nW;
  int ws, wt, ps, stride0, strideQ, dilation, ws_interior;
  int reflect_bounds, full_ws, patch_offset, off_Hq, off_Wq, dist_type, st_offset;
  float stride1;
} nls_args;

/* patch distance, integer centres: lib/csrc/search/nls_int.cu:4-83 (pt == 1) */
static float dist_int(const float* v0, const float* v1, const nls_args* a,
                      const int* ref_patch, const int* prop_patch){
  float dist = 0;
  int T=a->T, F=a->F, qH=a->qH, qW=a->qW, kH=a->kH, kW=a->kW;
  int r0 = bounds_i(ref_patch[0],T), p0 = bounds_i(prop_patch[0],T);
  int vr0 = inb_i(r0,T), vp0 = inb_i(p0,T);
  for (int pi=0; pi<a->ps; pi++){
    int r1 = ref_patch[1]+a->off_Hq+a->dilation*(pi+a->patch_offset);
    if (a->reflect_bounds) r1 = bounds_i(r1,qH);
    int p1 = prop_patch[1]+a->dilation*(pi+a->patch_offset);
    if (a->reflect_bounds) p1 = bounds_i(p1,kH);
    for (int pj=0; pj<a->ps; pj++){
      int r2 = ref_patch[2]+a->off_Wq+a->dilation*(pj+a->patch_offset);
      if (a->reflect_bounds) r2 = bounds_i(r2,qW);
      int p2 = prop_patch[2]+a->dilation*(pj+a->patch_offset);
      if (a->reflect_bounds) p2 = bounds_i(p2,kW);
      int valid = vr0 && inb_i(r1,qH) && inb_i(r2,qW) && vp0 && inb_i(p1,kH) && inb_i(p2,kW);
      if (!valid) continue;
      for (int c=0; c<F; c++){
        float x0 = v0[(((int64_t)r0*F+c)*qH+r1)*qW+r2];
        float x1 = v1[(((int64_t)p0*F+c)*kH+p1)*kW+p2];
        if (a->dist_type == 0) dist = fmaf(x0,x1,dist);
        else { float d = x0-x1; dist = fmaf(d,d,dist); }
      }
    }
  }
  return dist;
}

/* patch distance, float centres + bilinear k read: lib/csrc/search/nls_bilin2d.cu:3-90 */
static float dist_bilin(const float* v0, const float* v1, const nls_args* a,
                        const int* ref_patch, const float* prop_patch){
  float dist = 0;
  int T=a->T, F=a->F, qH=a->qH, qW=a->qW, kH=a->kH, kW=a->kW;
  int r0 = bounds_i(ref_patch[0],T);
  float p0 = bounds_f(prop_patch[0],T);
  int vr0 = inb_i(r0,T), vp0 = inb_f(p0,T);
  for (int pi=0; pi<a->ps; pi++){
    int r1 = ref_patch[1]+a->off_Hq+a->dilation*(pi+a->patch_offset);
    if (a->reflect_bounds) r1 = bounds_i(r1,qH);
    float p1 = prop_patch[1]+(float)(a->dilation*(pi+a->patch_offset));
    if (a->reflect_bounds) p1 = bounds_f(p1,kH);
    for (int pj=0; pj<a->ps; pj++){
      int r2 = ref_patch[2]+a->off_Wq+a->dilation*(pj+a->patch_offset);
      if (a->reflect_bounds) r2 = bounds_i(r2,qW);
      float p2 = prop_patch[2]+(float)(a->dilation*(pj+a->patch_offset));
      if (a->reflect_bounds) p2 = bounds_f(p2,kW);
      int valid = vr0 && inb_i(r1,qH) && inb_i(r2,qW) && vp0 && inb_f(p1,kH) && inb_f(p2,kW);
      if (!valid) continue;
      int pt0 = (int)rintf(p0);
      for (int c=0; c<F; c++){
        float x0 = v0[(((int64_t)r0*F+c)*qH+r1)*qW+r2];
        float x1 = bilin2d_interp(v1 + ((int64_t)pt0*F+c)*kH*kW, p1, p2, kH, kW);
        if (a->dist_type == 0) dist = fmaf(x0,x1,dist);
        else { float d = x0-x1; dist = fmaf(d,d,dist); }
      }
    }
  }
  return dist;
}

/*
 * Exhaustive forward search, every candidate materialised (what the reference kernel writes).
 *   int  : lib/csrc/search/non_local_search_int_kernel.cu:19-185
 *   float: lib/csrc/search/non_local_search_bilin2d_kernel.cu:20-193
 * flows: int32 (itype_int) or float32, shape (B,HD_f,T,W_t-st_offset,2,nH,nW)
 * dists: (B,HD,Q,W_t,ws,ws) ; inds: (...,3) int32 or float32
 */
void oracle_nls_forward(const float* vid0, const float* vid1, const void* flows_v,
                        float* dists, void* inds_v, const nls_args* a, int itype_int){
  int B=a->B, HD=a->HD, T=a->T, F=a->F, nH=a->nH, nW=a->nW, nHW=nH*nW, Q=T*nHW;
  int W_t = 2*a->wt+1, ws_og=a->ws, nF = W_t - a->st_offset;
  const int* flows_i = (const int*)flows_v; const float* flows_f = (const float*)flows_v;
  int* inds_i = (int*)inds_v; float* inds_f = (float*)inds_v;
  float invalid = (a->dist_type==0) ? -INFINITY : INFINITY;
  int64_t vq = (int64_t)T*F*a->qH*a->qW, vk = (int64_t)T*F*a->kH*a->kW;
  #pragma omp parallel for collapse(3) schedule(static)
  for (int b=0; b<B; b++) for (int hd=0; hd<HD; hd++) for (int qi=0; qi<Q; qi++){
    const float* v0 = vid0 + ((int64_t)b*HD+hd)*vq;
    const float* v1 = vid1 + ((int64_t)b*HD+hd)*vk;
    int hd_f = hd % a->HD_f;
    int ref_patch[3], adj_patch[3];
    if (itype_int){
      pixel_loc(ref_patch,qi,a->strideQ,nW,nHW,a->qH,a->qW);
      pixel_loc(adj_patch,qi,a->stride0,nW,nHW,a->kH,a->kW);
    }else{
      pixel_loc(ref_patch,qi,a->stride0,nW,nHW,a->qH,a->qW);
      memcpy(adj_patch,ref_patch,sizeof(ref_patch));
    }
    int sdiv = itype_int ? a->strideQ : a->stride0;
    int n_hi = ref_patch[1]/sdiv, n_wi = ref_patch[2]/sdiv;
    int ws = ws_og;
    if (itype_int){
      int btm_right = (n_hi == nH-1) || (n_wi == nW-1);
      ws = btm_right ? ws_og : a->ws_interior;
    }
    int t_max = time_tmax(ref_patch[0],T,a->wt);
    for (int st_i=0; st_i<W_t; st_i++){
      int t_next = adj_patch[0] + st_i;
      t_next = (t_next > t_max) ? t_max - st_i : t_next;
      int64_t fbase = 0;
      if (st_i >= a->st_offset)
        fbase = ((((int64_t)b*a->HD_f+hd_f)*T+adj_patch[0])*nF+(st_i-a->st_offset))*2;
      if (itype_int){
        int ah, aw;
        if (st_i >= a->st_offset){
          ah = adj_patch[1] + flows_i[((fbase+1)*nH+n_hi)*nW+n_wi];
          aw = adj_patch[2] + flows_i[((fbase+0)*nH+n_hi)*nW+n_wi];
          ah = bounds_i(ah,a->kH); aw = bounds_i(aw,a->kW);
        }else{ ah = adj_patch[1]; aw = adj_patch[2]; }
        int s1 = (int)a->stride1;
        int wsHalf = (ws-1)/2;
        int offh = search_off_i(ah,s1,wsHalf,ws,a->kH,a->full_ws);
        int offw = search_off_i(aw,s1,wsHalf,ws,a->kW,a->full_ws);
        for (int wi=0; wi<ws; wi++) for (int wj=0; wj<ws; wj++){
          int prop[3] = {t_next, ah + s1*(wi-offh), aw + s1*(wj-offw)};
          int valid = inb_i(prop[0],T) && inb_i(prop[1],a->kH) && inb_i(prop[2],a->kW);
          float d = valid ? dist_int(v0,v1,a,ref_patch,prop) : invalid;
          int64_t o = IDX6(b,hd,qi,st_i,wi,wj,HD,Q,W_t,ws_og,ws_og);
          dists[o] = d;
          inds_i[o*3+0] = prop[0]-adj_patch[0];
          inds_i[o*3+1] = prop[1]-adj_patch[1];
          inds_i[o*3+2] = prop[2]-adj_patch[2];
        }
      }else{
        float ah, aw;
        if (st_i >= a->st_offset){
          ah = ref_patch[1] + flows_f[((fbase+1)*nH+n_hi)*nW+n_wi];
          aw = ref_patch[2] + flows_f[((fbase+0)*nH+n_hi)*nW+n_wi];
          ah = bounds_f(ah,a->kH); aw = bounds_f(aw,a->kW);
        }else{ ah = (float)ref_patch[1]; aw = (float)ref_patch[2]; }
        float s1 = a->stride1;
        float wsHalf = (float)((ws-1)/2);
        float offh = search_off_f(ah,s1,wsHalf,ws,a->kH,a->full_ws);
        float offw = search_off_f(aw,s1,wsHalf,ws,a->kW,a->full_ws);
        for (int wi=0; wi<ws; wi++) for (int wj=0; wj<ws; wj++){
          float prop[3] = {(float)t_next, fmaf(s1,(float)wi-offh,ah), fmaf(s1,(float)wj-offw,aw)};
          int valid = inb_f(prop[0],T) && inb_f(prop[1],a->kH) && inb_f(prop[2],a->kW);
          float d = valid ? dist_bilin(v0,v1,a,ref_patch,prop) : invalid;
          int64_t o = IDX6(b,hd,qi,st_i,wi,wj,HD,Q,W_t,ws_og,ws_og);
          dists[o] = d;
          inds_f[o*3+0] = prop[0]-ref_patch[0];
          inds_f[o*3+1] = prop[1]-ref_patch[1];
          inds_f[o*3+2] = prop[2]-ref_patch[2];
        }
      }
    }
  }
}

/*
 * Refinement forward (every candidate materialised).
 *   int  : lib/csrc/search/refinement_int_kernel.cu:16-178  (incl. the flag permutation quirk: the
 *          launcher passes (restrict_radius, reflect_bounds, full_ws) into kernel parameters declared
 *          (reflect_bounds, full_ws, restrict_radius) -> the CALLER of this function applies it)
 *   float: lib/csrc/search/refinement_bilin2d_kernel.cu:18-185
 * flows (B,HD_f,T,nH,nW,Ks,3); dists (B,HD,T,nH,nW,Ks,wr,wr) pre-filled by the caller;
 * kselect (same shape as dists, int32), reflect (B,HD_f,T,nH,nW,Ks,2) uint8 (float path only).
 */
void oracle_refine_forward(const float* vid0, const float* vid1, const void* flows_v,
                           float* dists, void* inds_v, int* kselect, uint8_t* reflect,
                           const nls_args* a, int Ks, int wr, int itype_int){
  int B=a->B, HD=a->HD, T=a->T, nH=a->nH, nW=a->nW, nHW=nH*nW, Q=T*nHW;
  const int* flows_i = (const int*)flows_v; const float* flows_f = (const float*)flows_v;
  int* inds_i = (int*)inds_v; float* inds_f = (float*)inds_v;
  float invalid = (a->dist_type==0) ? -INFINITY : INFINITY;
  int64_t vq = (int64_t)T*a->F*a->qH*a->qW, vk = (int64_t)T*a->F*a->kH*a->kW;
  #pragma omp parallel for collapse(3) schedule(static)
  for (int b=0; b<B; b++) for (int hd=0; hd<HD; hd++) for (int qi=0; qi<Q; qi++){
    const float* v0 = vid0 + ((int64_t)b*HD+hd)*vq;
    const float* v1 = vid1 + ((int64_t)b*HD+hd)*vk;
    int hd_f = hd % a->HD_f;
    int ref_patch[3], adj_patch[3];
    if (itype_int){
      pixel_loc(ref_patch,qi,a->strideQ,nW,nHW,a->qH,a->qW);
      pixel_loc(adj_patch,qi,a->stride0,nW,nHW,a->kH,a->kW);
    }else{
      pixel_loc(ref_patch,qi,a->stride0,nW,nHW,a->qH,a->qW);
      memcpy(adj_patch,ref_patch,sizeof(ref_patch));
    }
    int ti = adj_patch[0], nh = adj_patch[1]/a->stride0, nw = adj_patch[2]/a->stride0;
    for (int ki=0; ki<Ks; ki++){
      int64_t fo = (((((int64_t)b*a->HD_f+hd_f)*T+ti)*nH+nh)*nW+nw)*Ks+ki;
      int64_t ob = (((((int64_t)b*HD+hd)*T+ti)*nH+nh)*nW+nw)*Ks+ki;
      if (itype_int){
        int f0=flows_i[fo*3], f1=flows_i[fo*3+1], f2=flows_i[fo*3+2];
        int pt0 = bounds_i(adj_patch[0]+f0,T);
        int ch = bounds_i(adj_patch[1]+f1,a->kH);
        int cw = bounds_i(adj_patch[2]+f2,a->kW);
        if (!((abs(f1) < 1e5) && (abs(f2) < 1e5))) continue;
        int s1 = (int)a->stride1;
        int wrHalf = (wr-1)/2;
        int offh = search_off_i(ch,s1,wrHalf,wr,a->kH,a->full_ws);
        int offw = search_off_i(cw,s1,wrHalf,wr,a->kW,a->full_ws);
        for (int wh=0; wh<wr; wh++) for (int ww=0; ww<wr; ww++){
          int prop[3] = {pt0, ch + s1*(wh-offh), cw + s1*(ww-offw)};
          int valid = inb_i(prop[0],T) && inb_i(prop[1],a->kH) && inb_i(prop[2],a->kW);
          float d = valid ? dist_int(v0,v1,a,ref_patch,prop) : invalid;
          int64_t o = (ob*wr+wh)*wr+ww;
          dists[o] = d;
          inds_i[o*3+0] = prop[0]-adj_patch[0];
          inds_i[o*3+1] = prop[1]-adj_patch[1];
          inds_i[o*3+2] = prop[2]-adj_patch[2];
        }
      }else{
        float f0=flows_f[fo*3], f1=flows_f[fo*3+1], f2=flows_f[fo*3+2];
        float pt0 = ref_patch[0] + floorf(f0+0.5f);
        float ch = ref_patch[1] + f1;
        float cw = ref_patch[2] + f2;
        pt0 = bounds_f(pt0,T);
        if (!((fabsf(f1) < 1e8) && (fabsf(f2) < 1e8))) continue;
        reflect[fo*2+0] = !inb_f(ch,a->kH);
        reflect[fo*2+1] = !inb_f(cw,a->kW);
        ch = bounds_f(ch,a->kH); cw = bounds_f(cw,a->kW);
        float s1 = a->stride1;
        float wrHalf = (float)((wr-1)/2);
        float offh = search_off_f(ch,s1,wrHalf,wr,a->kH,a->full_ws);
        float offw = search_off_f(cw,s1,wrHalf,wr,a->kW,a->full_ws);
        for (int wh=0; wh<wr; wh++) for (int ww=0; ww<wr; ww++){
          float prop[3] = {pt0, fmaf(s1,(float)wh-offh,ch), fmaf(s1,(float)ww-offw,cw)};
          int valid = inb_f(prop[0],T) && inb_f(prop[1],a->kH) && inb_f(prop[2],a->kW);
          float d = valid ? dist_bilin(v0,v1,a,ref_patch,prop) : invalid;
          int64_t o = (ob*wr+wh)*wr+ww;
          dists[o] = d;
          inds_f[o*3+0] = prop[0]-ref_patch[0];
          inds_f[o*3+1] = prop[1]-ref_patch[1];
          inds_f[o*3+2] = prop[2]-ref_patch[2];
          kselect[o] = ki;
        }
      }
    }
  }
}

/* lib/csrc/nn/anchor_self_kernel.cu:24-116 ; dists (N,L), inds (N,L,3), N = B*HD*Q */
void oracle_anchor_self(float* dists, void* inds_v, int64_t N, int L, int itype_int){
  int* inds_i = (int*)inds_v; float* inds_f = (float*)inds_v;
  #pragma omp parallel for schedule(static)
  for (int64_t n=0; n<N; n++){
    int min_idx = 0; float dmin = 10000;
    for (int s=0; s<L; s++){
      int64_t o = (n*L+s)*3;
      if (itype_int){
        if (inds_i[o]==0 && inds_i[o+1]==0 && inds_i[o+2]==0){ min_idx = s; break; }
      }else{
        float delta = 0;
        for (int ix=0; ix<3; ix++) delta += fabsf(inds_f[o+ix]);
        if (delta < 1e-8){ min_idx = s; break; }
        else if (delta < dmin){ min_idx = s; dmin = delta; }
      }
    }
    float tmp = dists[n*L]; dists[n*L] = dists[n*L+min_idx]; dists[n*L+min_idx] = tmp;
    int64_t o0 = n*L*3, o1 = (n*L+min_idx)*3;
    if (itype_int){
      int t[3] = {inds_i[o0],inds_i[o0+1],inds_i[o0+2]};
      inds_i[o0]=inds_i[o0+1]=inds_i[o0+2]=0;
      for (int ix=0; ix<3; ix++) inds_i[o1+ix]=t[ix];
    }else{
      float t[3] = {inds_f[o0],inds_f[o0+1],inds_f[o0+2]};
      inds_f[o0]=inds_f[o0+1]=inds_f[o0+2]=0;
      for (int ix=0; ix<3; ix++) inds_f[o1+ix]=t[ix];
    }
  }
}

/* lib/csrc/nn/anchor_self_kernel.cu:547-650 ; dists (B,HD,Q,G,K), inds (...,3), flows (B,HD_f,Q,G,3) */
void oracle_anchor_self_refine(float* dists, void* inds_v, const void* flows_v,
                               int B, int HD, int HD_f, int Q, int G, int K,
                               int stride0, int qH, int qW, int kH, int kW, int itype_int){
  int* inds_i = (int*)inds_v; float* inds_f = (float*)inds_v;
  const int* flows_i = (const int*)flows_v; const float* flows_f = (const float*)flows_v;
  int nH = (kH-1)/stride0+1, nW = (kW-1)/stride0+1, nHW = nH*nW;
  #pragma omp parallel for collapse(3) schedule(static)
  for (int b=0; b<B; b++) for (int hd=0; hd<HD; hd++) for (int qi=0; qi<Q; qi++){
    int hd_f = hd % HD_f;
    int iloc[3]; pixel_loc(iloc,qi,stride0,nW,nHW,qH,qW);
    for (int gi=0; gi<G; gi++){
      int64_t fo = ((((int64_t)b*HD_f+hd_f)*Q+qi)*G+gi)*3;
      int64_t base = ((((int64_t)b*HD+hd)*Q+qi)*G+gi)*K;
      int min_idx = 0; float dmin = 10000;
      if (itype_int){
        int loc[3];
        loc[0] = (int)round((double)flows_i[fo]);
        loc[1] = bounds_i(iloc[1]+flows_i[fo+1],kH)-iloc[1];
        loc[2] = bounds_i(iloc[2]+flows_i[fo+2],kW)-iloc[2];
        for (int s=0; s<K; s++){
          int64_t o=(base+s)*3;
          if (inds_i[o]==loc[0] && inds_i[o+1]==loc[1] && inds_i[o+2]==loc[2]){ min_idx=s; break; }
        }
        float tmp=dists[base]; dists[base]=dists[base+min_idx]; dists[base+min_idx]=tmp;
        int64_t o0=base*3, o1=(base+min_idx)*3;
        int t[3]={inds_i[o0],inds_i[o0+1],inds_i[o0+2]};
        for (int ix=0; ix<3; ix++) inds_i[o0+ix]=loc[ix];
        for (int ix=0; ix<3; ix++) inds_i[o1+ix]=t[ix];
      }else{
        float loc[3];
        loc[0] = roundf(flows_f[fo]);
        loc[1] = bounds_f(iloc[1]+flows_f[fo+1],kH)-iloc[1];
        loc[2] = bounds_f(iloc[2]+flows_f[fo+2],kW)-iloc[2];
        for (int s=0; s<K; s++){
          int64_t o=(base+s)*3; float delta=0;
          for (int ix=0; ix<3; ix++) delta += fabsf(inds_f[o+ix]-loc[ix]);
          if (delta < 1e-8){ min_idx=s; break; }
          else if (delta < dmin){ min_idx=s; dmin=delta; }
        }
        float tmp=dists[base]; dists[base]=dists[base+min_idx]; dists[base+min_idx]=tmp;
        int64_t o0=base*3, o1=(base+min_idx)*3;
        float t[3]={inds_f[o0],inds_f[o0+1],inds_f[o0+2]};
        for (int ix=0; ix<3; ix++) inds_f[o0+ix]=loc[ix];
        for (int ix=0; ix<3; ix++) inds_f[o1+ix]=t[ix];
      }
    }
  }
}

/* lib/csrc/nn/anchor_self_kernel.cu:187-307 ; dists (B,HD,Q,G,S), inds (...,3), flows (B,HD_f,T,nflow,2,nH,nW),
 * Q = T*nH*nW; G = 2wt+1 or 2wt frame groups; st_offset = G - nflow (anchor_self_kernel.cu:327) */
void oracle_anchor_self_time(float* dists, void* inds_v, const void* flows_v,
                             int B, int HD, int HD_f, int T, int G, int S, int nflow,
                             int wt, int stride0, int qH, int qW, int kH, int kW, int itype_int){
  int* inds_i = (int*)inds_v; float* inds_f = (float*)inds_v;
  const int* flows_i = (const int*)flows_v; const float* flows_f = (const float*)flows_v;
  int nH = (kH-1)/stride0+1, nW = (kW-1)/stride0+1, nHW = nH*nW;
  int Q = T*nHW, st_offset = G - nflow;
  #pragma omp parallel for collapse(3) schedule(static)
  for (int b=0; b<B; b++) for (int hd=0; hd<HD; hd++) for (int qi=0; qi<Q; qi++){
    int hd_f = hd % HD_f;
    int iloc[3]; pixel_loc(iloc,qi,stride0,nW,nHW,qH,qW);
    int n_hi = iloc[1]/stride0, n_wi = iloc[2]/stride0;
    int t_max = time_tmax(iloc[0],T,wt);
    for (int st=0; st<G; st++){
      int t_next = iloc[0] + st;
      t_next = (t_next > t_max) ? t_max - st : t_next;
      int64_t base = ((((int64_t)b*HD+hd)*Q+qi)*G+st)*S;
      int64_t fo = 0;
      if (st >= st_offset)
        fo = (((((int64_t)b*HD_f+hd_f)*T+iloc[0])*nflow+(st-st_offset))*2)*(int64_t)nHW + (int64_t)n_hi*nW + n_wi;
      int min_idx = 0; float dmin = 10000;
      if (itype_int){
        int loc[3] = {t_next - iloc[0], 0, 0};
        if (st >= st_offset){
          loc[1] = bounds_i(iloc[1]+flows_i[fo+nHW],kH)-iloc[1];
          loc[2] = bounds_i(iloc[2]+flows_i[fo],kW)-iloc[2];
        }
        for (int s=0; s<S; s++){
          int64_t o=(base+s)*3;
          if (inds_i[o]==loc[0] && inds_i[o+1]==loc[1] && inds_i[o+2]==loc[2]){ min_idx=s; break; }
        }
        float tmp=dists[base]; dists[base]=dists[base+min_idx]; dists[base+min_idx]=tmp;
        int64_t o0=base*3, o1=(base+min_idx)*3;
        int t[3]={inds_i[o0],inds_i[o0+1],inds_i[o0+2]};
        for (int ix=0; ix<3; ix++) inds_i[o0+ix]=loc[ix];
        for (int ix=0; ix<3; ix++) inds_i[o1+ix]=t[ix];
      }else{
        float loc[3] = {(float)(t_next - iloc[0]), 0.f, 0.f};
        if (st >= st_offset){
          loc[1] = bounds_f(iloc[1]+flows_f[fo+nHW],kH)-iloc[1];
          loc[2] = bounds_f(iloc[2]+flows_f[fo],kW)-iloc[2];
        }
        for (int s=0; s<S; s++){
          int64_t o=(base+s)*3; float delta=0;
          for (int ix=0; ix<3; ix++) delta += fabsf(inds_f[o+ix]-loc[ix]);
          if (delta < 1e-8){ min_idx=s; break; }
          else if (delta < dmin){ min_idx=s; dmin=delta; }
        }
        float tmp=dists[base]; dists[base]=dists[base+min_idx]; dists[base+min_idx]=tmp;
        int64_t o0=base*3, o1=(base+min_idx)*3;
        float t[3]={inds_f[o0],inds_f[o0+1],inds_f[o0+2]};
        for (int ix=0; ix<3; ix++) inds_f[o0+ix]=loc[ix];
        for (int ix=0; ix<3; ix++) inds_f[o1+ix]=t[ix];
      }
    }
  }
}

/*
 * Search backward.  mode 0: int   (non_local_search_int_kernel.cu:273-355 + nls_int.cu:88-167)
 *                   mode 1: bilin (non_local_search_bilin2d_kernel.cu:285-371 + nls_bilin2d.cu:95-186)
 *                   mode 2: bilin + flows grad, NLS flows layout
 *                           (non_local_search_bilin2d_kernel.cu:464-598 + nls_bilin2d.cu:189-291)
 *                   mode 3: bilin + flows grad, refine layout routed through kselect/reflect
 *                           (refinement_bilin2d_kernel.cu:359-491, shared_kernel.cu:1001-1026)
 * grad_dists (B,HD,T,nH,nW,K), inds (...,K,3); grads pre-zeroed by the caller.
 */
void oracle_search_backward(float* gvid0, float* gvid1, float* gflows,
                            const float* vid0, const float* vid1, const float* flows,
                            const float* grad_dists, const float* grad_inds, const void* inds_v,
                            const int* kselect, const uint8_t* reflect,
                            const nls_args* a, int K, int Ks, int mode){
  int B=a->B, HD=a->HD, T=a->T, F=a->F, nH=a->nH, nW=a->nW, nHW=nH*nW, Q=T*nHW;
  int qH=a->qH, qW=a->qW, kH=a->kH, kW=a->kW;
  const int* inds_i = (const int*)inds_v; const float* inds_f = (const float*)inds_v;
  int64_t vq = (int64_t)T*F*qH*qW, vk = (int64_t)T*F*kH*kW;
  int W_t = 2*a->wt+1, nF = W_t - a->st_offset;
  #pragma omp parallel for collapse(3) schedule(static)
  for (int b=0; b<B; b++) for (int hd=0; hd<HD; hd++) for (int qi=0; qi<Q; qi++){
    const float* v0 = vid0 + ((int64_t)b*HD+hd)*vq; float* g0 = gvid0 + ((int64_t)b*HD+hd)*vq;
    const float* v1 = vid1 + ((int64_t)b*HD+hd)*vk; float* g1 = gvid1 + ((int64_t)b*HD+hd)*vk;
    int hd_f = hd % a->HD_f;
    int ref_patch[3], adj_patch[3];
    if (mode == 0){
      pixel_loc(ref_patch,qi,a->strideQ,nW,nHW,qH,qW);
      pixel_loc(adj_patch,qi,a->stride0,nW,nHW,kH,kW);
    }else{
      pixel_loc(ref_patch,qi,a->stride0,nW,nHW,qH,qW);
      memcpy(adj_patch,ref_patch,sizeof(ref_patch));
    }
    int sdiv = (mode==0) ? a->strideQ : a->stride0;
    int ti = ref_patch[0], nh = ref_patch[1]/sdiv, nw = ref_patch[2]/sdiv;
    for (int ki=0; ki<K; ki++){
      int64_t o = (((((int64_t)b*HD+hd)*T+ti)*nH+nh)*nW+nw)*K+ki;
      float weight = grad_dists[o];
      if (mode == 0){
        int prop_patch[3] = {adj_patch[0]+inds_i[o*3], adj_patch[1]+inds_i[o*3+1], adj_patch[2]+inds_i[o*3+2]};
        int r0 = bounds_i(ref_patch[0],T), p0 = bounds_i(prop_patch[0],T);
        int vr0 = inb_i(r0,T), vp0 = inb_i(p0,T);
        for (int pi=0; pi<a->ps; pi++){
          int r1 = ref_patch[1]+a->off_Hq+a->dilation*(pi+a->patch_offset);
          if (a->reflect_bounds) r1 = bounds_i(r1,qH);
          int p1 = prop_patch[1]+a->dilation*(pi+a->patch_offset);
          if (a->reflect_bounds) p1 = bounds_i(p1,kH);
          for (int pj=0; pj<a->ps; pj++){
            int r2 = ref_patch[2]+a->off_Wq+a->dilation*(pj+a->patch_offset);
            if (a->reflect_bounds) r2 = bounds_i(r2,qW);
            int p2 = prop_patch[2]+a->dilation*(pj+a->patch_offset);
            if (a->reflect_bounds) p2 = bounds_i(p2,kW);
            int valid = vr0 && inb_i(r1,qH) && inb_i(r2,qW) && vp0 && inb_i(p1,kH) && inb_i(p2,kW);
            if (!valid) continue;
            for (int c=0; c<F; c++){
              int64_t i0 = (((int64_t)r0*F+c)*qH+r1)*qW+r2;
              int64_t i1 = (((int64_t)p0*F+c)*kH+p1)*kW+p2;
              float a0, a1;
              if (a->dist_type == 0){ a0 = weight*v1[i1]; a1 = weight*v0[i0]; }
              else { float pix = 2*weight*(v0[i0]-v1[i1]); a0 = pix; a1 = -pix; }
              #pragma omp atomic
              g0[i0] += a0;
              #pragma omp atomic
              g1[i1] += a1;
            }
          }
        }
        continue;
      }
      /* float modes */
      float prop_patch[3] = {ref_patch[0]+inds_f[o*3], ref_patch[1]+inds_f[o*3+1], ref_patch[2]+inds_f[o*3+2]};
      int signH_c=1, signW_c=1, si=-1, kj=0;
      float iw0=0, iw1=0;
      if (mode >= 2){ iw0 = grad_inds[o*3+1]; iw1 = grad_inds[o*3+2]; }
      if (mode == 3){
        kj = kselect[o];
        int64_t ro = ((((((int64_t)b*a->HD_f+hd_f)*T+ti)*nH+nh)*nW+nw)*Ks+kj);
        signH_c = reflect[ro*2+0] ? -1 : 1;
        signW_c = reflect[ro*2+1] ? -1 : 1;
        #pragma omp atomic
        gflows[ro*3+1] += signH_c*iw0;
        #pragma omp atomic
        gflows[ro*3+2] += signW_c*iw1;
      }
      if (mode >= 2){
        int valid_pp = inb_f(prop_patch[0],T) && inb_f(prop_patch[1],kH) && inb_f(prop_patch[2],kW);
        if (!valid_pp) continue;
      }
      if (mode == 2){
        int t_max = time_tmax(ti,T,a->wt);
        int dt = (int)prop_patch[0] - ti;
        int dto = t_max - ti;
        si = (dt >= 0) ? (dt - a->st_offset) : (dto - dt - a->st_offset);
      }
      float acc[8]; for (int i=0;i<8;i++) acc[i]=0;
      int r0 = bounds_i(ref_patch[0],T);
      float p0 = bounds_f(prop_patch[0],T);
      int vr0 = inb_i(r0,T), vp0 = inb_f(p0,T);
      for (int pi=0; pi<a->ps; pi++){
        int r1 = ref_patch[1]+a->off_Hq+a->dilation*(pi+a->patch_offset);
        if (a->reflect_bounds) r1 = bounds_i(r1,qH);
        float p1 = prop_patch[1]+(float)(a->dilation*(pi+a->patch_offset));
        int signH = inb_f(p1,kH) ? 1 : -1;
        if (a->reflect_bounds) p1 = bounds_f(p1,kH);
        for (int pj=0; pj<a->ps; pj++){
          int r2 = ref_patch[2]+a->off_Wq+a->dilation*(pj+a->patch_offset);
          if (a->reflect_bounds) r2 = bounds_i(r2,qW);
          float p2 = prop_patch[2]+(float)(a->dilation*(pj+a->patch_offset));
          int signW = inb_f(p2,kW) ? 1 : -1;
          if (a->reflect_bounds) p2 = bounds_f(p2,kW);
          int valid = vr0 && inb_i(r1,qH) && inb_i(r2,qW) && vp0 && inb_f(p1,kH) && inb_f(p2,kW);
          if (!valid) continue;
          int pt0 = (int)rintf(p0);
          for (int c=0; c<F; c++){
            int64_t i0 = (((int64_t)r0*F+c)*qH+r1)*qW+r2;
            const float* plane = v1 + ((int64_t)pt0*F+c)*kH*kW;
            float pix0 = v0[i0];
            float pix1 = bilin2d_interp(plane,p1,p2,kH,kW);
            float dD;
            if (a->dist_type == 0) dD = weight*pix1; else dD = weight*2*(pix0-pix1);
            #pragma omp atomic
            g0[i0] += dD;
            if (a->dist_type == 0) dD = weight*pix0; else dD = -dD;
            bilin2d_assign(dD,p1,p2,kH,kW,g1 + ((int64_t)pt0*F+c)*kH*kW);
            if (mode >= 2) update_dflows(acc,dD,p1,p2,kH,kW,signH,signW,plane);
          }
        }
      }
      if (mode == 2){
        if (si < 0) continue;
        int64_t fb = ((((int64_t)b*a->HD_f+hd_f)*T+ti)*nF+si)*2;
        float hi = ref_patch[1] + flows[((fb+1)*nH+nh)*nW+nw];
        float wi = ref_patch[2] + flows[((fb+0)*nH+nh)*nW+nw];
        int sH = inb_f(hi,kH) ? 1 : -1, sW = inb_f(wi,kW) ? 1 : -1;
        float* gw = gflows + ((fb+0)*nH+nh)*nW+nw;
        float* gh = gflows + ((fb+1)*nH+nh)*nW+nw;
        for (int ix=0; ix<2; ix++) for (int jx=0; jx<2; jx++){
          int kx = ix*4+jx*2;
          float aw = sW*acc[kx], ah = sH*acc[kx+1];
          #pragma omp atomic
          *gw += aw;
          #pragma omp atomic
          *gh += ah;
        }
        float bw = sW*iw1, bh = sH*iw0;
        #pragma omp atomic
        *gw += bw;
        #pragma omp atomic
        *gh += bh;
      }else if (mode == 3){
        int64_t ro = ((((((int64_t)b*a->HD_f+hd_f)*T+ti)*nH+nh)*nW+nw)*Ks+kj);
        for (int ix=0; ix<2; ix++) for (int jx=0; jx<2; jx++){
          int kx = ix*4+jx*2;
          float aw = signW_c*acc[kx], ah = signH_c*acc[kx+1];
          #pragma omp atomic
          gflows[ro*3+2] += aw;
          #pragma omp atomic
          gflows[ro*3+1] += ah;
        }
      }
    }
  }
}

/*
 * Weighted patch sum == gather_add forward: lib/csrc/agg/gather_add_kernel.cu:15-136.
 * vid (B,HD,T,F,H,W); weights (B,HD,Q,K); flows (B,HD,Q,K,3) int32/float32; out pre-zeroed,
 * counts (H,W) int32 pre-zeroed.  strideIn == strideOut, out size == in size (reference quirk A.9-3).
 */
void oracle_gather_add_forward(float* out, int* counts, const float* vid, const float* weights,
                               const void* flows_v, int B, int HD, int T, int F, int H, int W,
                               int Q, int K, int ps, int stride0, int dilation, int reflect_bounds,
                               int patch_offset, int itype_int){
  const int* flows_i = (const int*)flows_v; const float* flows_f = (const float*)flows_v;
  int nW = (W-1)/stride0+1, nHW = nW*((H-1)/stride0+1);
  int64_t vsz = (int64_t)T*F*H*W;
  #pragma omp parallel for collapse(3) schedule(static)
  for (int b=0; b<B; b++) for (int hd=0; hd<HD; hd++) for (int qi=0; qi<Q; qi++){
    const float* v = vid + ((int64_t)b*HD+hd)*vsz; float* o = out + ((int64_t)b*HD+hd)*vsz;
    int ref[3]; pixel_loc(ref,qi,stride0,nW,nHW,H,W);
    for (int ki=0; ki<K; ki++){
      int64_t fo = ((((int64_t)b*HD+hd)*Q+qi)*K+ki);
      float weight = weights[fo];
      if (itype_int){
        int nl[3] = {ref[0]+flows_i[fo*3], ref[1]+flows_i[fo*3+1], ref[2]+flows_i[fo*3+2]};
        if (!((abs(nl[1]) < 1e7) && (abs(nl[2]) < 1e7))) continue;
        nl[0]=bounds_i(nl[0],T); nl[1]=bounds_i(nl[1],H); nl[2]=bounds_i(nl[2],W);
        for (int pi=0; pi<ps; pi++) for (int pj=0; pj<ps; pj++){
          int r1 = ref[1]+dilation*(pi+patch_offset), r2 = ref[2]+dilation*(pj+patch_offset);
          if (!(inb_i(ref[0],T) && inb_i(r1,H) && inb_i(r2,W))) continue;
          if (ref[0]==0 && b==0 && hd==0 && ki==0){
            #pragma omp atomic
            counts[r1*W+r2] += 1;
          }
          int n1 = nl[1]+dilation*(pi+patch_offset); if (reflect_bounds) n1 = bounds_i(n1,H);
          int n2 = nl[2]+dilation*(pj+patch_offset); if (reflect_bounds) n2 = bounds_i(n2,W);
          if (!(inb_i(nl[0],T) && inb_i(n1,H) && inb_i(n2,W))) continue;
          int n0 = reflect_bounds ? bounds_i(nl[0],T) : nl[0];
          if (!(n0 >= 0 && n0 < T)) continue;
          for (int c=0; c<F; c++){
            float add = weight*v[(((int64_t)n0*F+c)*H+n1)*W+n2];
            #pragma omp atomic
            o[(((int64_t)ref[0]*F+c)*H+r1)*W+r2] += add;
          }
        }
      }else{
        float nl[3] = {ref[0]+flows_f[fo*3], ref[1]+flows_f[fo*3+1], ref[2]+flows_f[fo*3+2]};
        if (!((fabsf(nl[1]) < 1e7) && (fabsf(nl[2]) < 1e7))) continue;
        nl[0]=bounds_f(nl[0],T); nl[1]=bounds_f(nl[1],H); nl[2]=bounds_f(nl[2],W);
        for (int pi=0; pi<ps; pi++) for (int pj=0; pj<ps; pj++){
          float r1 = ref[1]+dilation*(pi+patch_offset), r2 = ref[2]+dilation*(pj+patch_offset);
          if (!(inb_f((float)ref[0],T) && inb_f(r1,H) && inb_f(r2,W))) continue;
          if (ref[0]==0 && b==0 && hd==0 && ki==0){
            #pragma omp atomic
            counts[(int)r1*W+(int)r2] += 1;
          }
          float n1 = nl[1]+(float)(dilation*(pi+patch_offset)); if (reflect_bounds) n1 = bounds_f(n1,H);
          float n2 = nl[2]+(float)(dilation*(pj+patch_offset)); if (reflect_bounds) n2 = bounds_f(n2,W);
          if (!(inb_f(nl[0],T) && inb_f(n1,H) && inb_f(n2,W))) continue;
          float n0 = reflect_bounds ? bounds_f(nl[0],T) : nl[0];
          if (!(n0 >= 0 && n0 < T)) continue;
          int n0i = (int)n0;
          for (int c=0; c<F; c++){
            float pix = bilin2d_interp(v + ((int64_t)n0i*F+c)*H*W, n1, n2, H, W);
            float add = weight*pix;
            #pragma omp atomic
            o[(((int64_t)ref[0]*F+c)*H+(int)r1)*W+(int)r2] += add;
          }
        }
      }
    }
  }
}

/*
 * gather_add backward.  int : lib/csrc/agg/gather_add_kernel.cu:201-319
 *                       float: lib/csrc/agg/gather_add_kernel.cu:369-514 + shared_kernel.cu:773-827
 * gout is the incoming gradient ALREADY divided by counts (lib/stnls/agg/gather_add.py:133-134).
 */
void oracle_gather_add_backward(float* gvid, float* gweights, float* gflows,
                                const float* gout, const float* vid, const float* weights,
                                const void* flows_v, int B, int HD, int T, int F, int H, int W,
                                int Q, int K, int ps, int stride0, int dilation, int reflect_bounds,
                                int patch_offset, int itype_int){
  const int* flows_i = (const int*)flows_v; const float* flows_f = (const float*)flows_v;
  int nW = (W-1)/stride0+1, nHW = nW*((H-1)/stride0+1);
  int64_t vsz = (int64_t)T*F*H*W;
  #pragma omp parallel for collapse(3) schedule(static)
  for (int b=0; b<B; b++) for (int hd=0; hd<HD; hd++) for (int qi=0; qi<Q; qi++){
    const float* v = vid + ((int64_t)b*HD+hd)*vsz; float* gv = gvid + ((int64_t)b*HD+hd)*vsz;
    const float* go = gout + ((int64_t)b*HD+hd)*vsz;
    int ref[3]; pixel_loc(ref,qi,stride0,nW,nHW,H,W);
    for (int ki=0; ki<K; ki++){
      int64_t fo = ((((int64_t)b*HD+hd)*Q+qi)*K+ki);
      float weight = weights[fo];
      if (itype_int){
        int nl[3] = {ref[0]+flows_i[fo*3], ref[1]+flows_i[fo*3+1], ref[2]+flows_i[fo*3+2]};
        if (!((abs(nl[1]) < 1e7) && (abs(nl[2]) < 1e7))) continue;
        nl[0]=bounds_i(nl[0],T); nl[1]=bounds_i(nl[1],H); nl[2]=bounds_i(nl[2],W);
        for (int pi=0; pi<ps; pi++) for (int pj=0; pj<ps; pj++){
          int r1 = ref[1]+dilation*(pi+patch_offset), r2 = ref[2]+dilation*(pj+patch_offset);
          if (!(inb_i(ref[0],T) && inb_i(r1,H) && inb_i(r2,W))) continue;
          int n1 = nl[1]+dilation*(pi+patch_offset); if (reflect_bounds) n1 = bounds_i(n1,H);
          int n2 = nl[2]+dilation*(pj+patch_offset); if (reflect_bounds) n2 = bounds_i(n2,W);
          if (!(inb_i(nl[0],T) && inb_i(n1,H) && inb_i(n2,W))) continue;
          int n0 = reflect_bounds ? bounds_i(nl[0],T) : nl[0];
          if (!(n0 >= 0 && n0 < T)) continue;
          float acc = 0;
          for (int c=0; c<F; c++){
            float grad = go[(((int64_t)ref[0]*F+c)*H+r1)*W+r2];
            int64_t iv = (((int64_t)n0*F+c)*H+n1)*W+n2;
            float add = weight*grad;
            #pragma omp atomic
            gv[iv] += add;
            acc += grad*v[iv];
          }
          gweights[fo] += acc;   /* (qi,ki) is owned by this iteration: no race */
        }
      }else{
        float nl[3] = {ref[0]+flows_f[fo*3], ref[1]+flows_f[fo*3+1], ref[2]+flows_f[fo*3+2]};
        if (!((fabsf(nl[1]) < 1e7) && (fabsf(nl[2]) < 1e7))) continue;
        nl[0]=bounds_f(nl[0],T);
        int sH0 = inb_f(nl[1],H) ? 1 : -1; nl[1]=bounds_f(nl[1],H);
        int sW0 = inb_f(nl[2],W) ? 1 : -1; nl[2]=bounds_f(nl[2],W);
        for (int pi=0; pi<ps; pi++) for (int pj=0; pj<ps; pj++){
          int r1 = ref[1]+dilation*(pi+patch_offset), r2 = ref[2]+dilation*(pj+patch_offset);
          if (!(inb_i(ref[0],T) && inb_i(r1,H) && inb_i(r2,W))) continue;
          float n1 = nl[1]+(float)(dilation*(pi+patch_offset));
          int signH = inb_f(n1,H) ? sH0 : -sH0;
          if (reflect_bounds) n1 = bounds_f(n1,H);
          float n2 = nl[2]+(float)(dilation*(pj+patch_offset));
          int signW = inb_f(n2,W) ? sW0 : -sW0;
          if (reflect_bounds) n2 = bounds_f(n2,W);
          if (!(inb_f(nl[0],T) && inb_f(n1,H) && inb_f(n2,W))) continue;
          int nti = reflect_bounds ? (int)bounds_f(nl[0],T) : (int)nl[0];
          float acc=0, accH=0, accW=0;
          for (int c=0; c<F; c++){
            float grad = go[(((int64_t)ref[0]*F+c)*H+r1)*W+r2];
            const float* plane = v + ((int64_t)nti*F+c)*H*W;
            float* gplane = gv + ((int64_t)nti*F+c)*H*W;
            float wg = weight*grad, pix=0, igW=0, igH=0;
            for (int ix=0; ix<2; ix++) for (int jx=0; jx<2; jx++){   /* shared_kernel.cu:773-827 */
              int h_i = (int)truncf(n1+ix), w_i = (int)truncf(n2+jx);
              float gH = fmaxf(0.f,1-fabsf(n1-h_i)), gW = fmaxf(0.f,1-fabsf(n2-w_i));
              float iwt = gH*gW;
              if (!inb_i(h_i,H)) continue;
              if (!inb_i(w_i,W)) continue;
              float vp = plane[(int64_t)h_i*W+w_i];
              pix += iwt*vp;
              igW += (jx==0) ? -gH*vp : gH*vp;
              igH += (ix==0) ? -gW*vp : gW*vp;
              float add = iwt*wg;
              #pragma omp atomic
              gplane[(int64_t)h_i*W+w_i] += add;
            }
            acc += grad*pix; accW += grad*igW; accH += grad*igH;
          }
          gweights[fo] += acc;
          gflows[fo*3+1] += weight*accH*signH;
          gflows[fo*3+2] += weight*accW*signW;
        }
      }
    }
  }
}

/*
 * NonLocalGather (K-stacked aggregation) forward.
 *   int  : lib/csrc/agg/gather_int_kernel.cu:19-97   + lib/csrc/agg/gather_int.cu:15-94
 *   float: lib/csrc/agg/gather_bilin2d_kernel.cu:24-98 + lib/csrc/agg/gather_bilin2d.cu:14-93
 * vid (B,HD,T,F,H,W); weights (B,HD,Q,K); inds (B,HD,Q,K,3); stack (B,HD,K,T,F,H,W) pre-zeroed;
 * counts (H,W) int32 pre-zeroed (the reference keeps one identical map per (b,hd)).  pt == 1.
 */
void oracle_gather_forward(float* stack, int* counts, const float* vid, const float* weights,
                           const void* inds_v, int B, int HD, int T, int F, int H, int W,
                           int Q, int K, int ps, int stride0, int dilation, int reflect_bounds,
                           int patch_offset, int itype_int){
  const int* inds_i = (const int*)inds_v; const float* inds_f = (const float*)inds_v;
  int nW = (W-1)/stride0+1, nHW = nW*((H-1)/stride0+1);
  int64_t vsz = (int64_t)T*F*H*W;
  #pragma omp parallel for collapse(3) schedule(static)
  for (int b=0; b<B; b++) for (int hd=0; hd<HD; hd++) for (int qi=0; qi<Q; qi++){
    const float* v = vid + ((int64_t)b*HD+hd)*vsz;
    int ref[3]; pixel_loc(ref,qi,stride0,nW,nHW,H,W);
    for (int ki=0; ki<K; ki++){
      int64_t fo = ((((int64_t)b*HD+hd)*Q+qi)*K+ki);
      float weight = weights[fo];
      float* o = stack + (((int64_t)b*HD+hd)*K+ki)*vsz;
      int r0 = bounds_i(ref[0],T);
      if (itype_int){
        int nl0 = bounds_i(ref[0]+inds_i[fo*3],T), nl1 = bounds_i(ref[1]+inds_i[fo*3+1],H),
            nl2 = bounds_i(ref[2]+inds_i[fo*3+2],W);
        int n0 = bounds_i(nl0,T);
        for (int pi=0; pi<ps; pi++) for (int pj=0; pj<ps; pj++){
          int r1 = ref[1]+dilation*(pi+patch_offset), r2 = ref[2]+dilation*(pj+patch_offset);
          int vref = inb_i(r0,T) && inb_i(r1,H) && inb_i(r2,W);
          int n1 = nl1+dilation*(pi+patch_offset); if (reflect_bounds) n1 = bounds_i(n1,H);
          int n2 = nl2+dilation*(pj+patch_offset); if (reflect_bounds) n2 = bounds_i(n2,W);
          int vnl = inb_i(n0,T) && inb_i(n1,H) && inb_i(n2,W);
          if (ki==0 && vref && r0==0 && b==0 && hd==0){
            #pragma omp atomic
            counts[r1*W+r2] += 1;
          }
          if (!(vref && vnl)) continue;
          for (int c=0; c<F; c++){
            float add = weight*v[(((int64_t)n0*F+c)*H+n1)*W+n2];
            #pragma omp atomic
            o[(((int64_t)r0*F+c)*H+r1)*W+r2] += add;
          }
        }
      }else{
        float nl0 = bounds_f(ref[0]+inds_f[fo*3],T), nl1 = bounds_f(ref[1]+inds_f[fo*3+1],H),
              nl2 = bounds_f(ref[2]+inds_f[fo*3+2],W);
        float n0 = bounds_f(nl0,T);
        for (int pi=0; pi<ps; pi++) for (int pj=0; pj<ps; pj++){
          int r1 = ref[1]+dilation*(pi+patch_offset), r2 = ref[2]+dilation*(pj+patch_offset);
          int vref = inb_i(r0,T) && inb_i(r1,H) && inb_i(r2,W);
          float n1 = nl1+(float)(dilation*(pi+patch_offset)); if (reflect_bounds) n1 = bounds_f(n1,H);
          float n2 = nl2+(float)(dilation*(pj+patch_offset)); if (reflect_bounds) n2 = bounds_f(n2,W);
          int vnl = inb_f(n0,T) && inb_f(n1,H) && inb_f(n2,W);
          if (ki==0 && vref && r0==0 && b==0 && hd==0){
            #pragma omp atomic
            counts[r1*W+r2] += 1;
          }
          if (!(vref && vnl)) continue;
          int ti = (int)lrintf(roundf(n0));
          for (int c=0; c<F; c++){
            float pix = bilin2d_interp(v + ((int64_t)ti*F+c)*H*W, n1, n2, H, W);
            float add = weight*pix;
            #pragma omp atomic
            o[(((int64_t)r0*F+c)*H+r1)*W+r2] += add;
          }
        }
      }
    }
  }
}

/*
 * NonLocalGather backward.
 *   int  : lib/csrc/agg/gather_int_kernel.cu:150-230    + lib/csrc/agg/gather_int.cu:98-174
 *   float: lib/csrc/agg/gather_bilin2d_kernel.cu:160-250 + lib/csrc/agg/gather_bilin2d.cu:95-236
 *          (taps outside the frame are REFLECTED here, :188-189, not dropped as in gather_add)
 * gstack (B,HD,K,T,F,H,W) is the incoming gradient ALREADY divided by counts (gather.py:231).
 */
void oracle_gather_backward(float* gvid, float* gweights, float* ginds,
                            const float* gstack, const float* vid, const float* weights,
                            const void* inds_v, int B, int HD, int T, int F, int H, int W,
                            int Q, int K, int ps, int stride0, int dilation, int reflect_bounds,
                            int patch_offset, int itype_int){
  const int* inds_i = (const int*)inds_v; const float* inds_f = (const float*)inds_v;
  int nW = (W-1)/stride0+1, nHW = nW*((H-1)/stride0+1);
  int64_t vsz = (int64_t)T*F*H*W;
  #pragma omp parallel for collapse(3) schedule(static)
  for (int b=0; b<B; b++) for (int hd=0; hd<HD; hd++) for (int qi=0; qi<Q; qi++){
    const float* v = vid + ((int64_t)b*HD+hd)*vsz; float* gv = gvid + ((int64_t)b*HD+hd)*vsz;
    int ref[3]; pixel_loc(ref,qi,stride0,nW,nHW,H,W);
    int r0 = bounds_i(ref[0],T);
    for (int ki=0; ki<K; ki++){
      int64_t fo = ((((int64_t)b*HD+hd)*Q+qi)*K+ki);
      float weight = weights[fo];
      const float* go = gstack + (((int64_t)b*HD+hd)*K+ki)*vsz;
      if (itype_int){
        int nl0 = bounds_i(ref[0]+inds_i[fo*3],T), nl1 = bounds_i(ref[1]+inds_i[fo*3+1],H),
            nl2 = bounds_i(ref[2]+inds_i[fo*3+2],W);
        int n0 = bounds_i(nl0,T);
        for (int pi=0; pi<ps; pi++) for (int pj=0; pj<ps; pj++){
          int r1 = ref[1]+dilation*(pi+patch_offset), r2 = ref[2]+dilation*(pj+patch_offset);
          int n1 = nl1+dilation*(pi+patch_offset); if (reflect_bounds) n1 = bounds_i(n1,H);
          int n2 = nl2+dilation*(pj+patch_offset); if (reflect_bounds) n2 = bounds_i(n2,W);
          if (!(inb_i(r0,T) && inb_i(r1,H) && inb_i(r2,W))) continue;
          if (!(inb_i(n0,T) && inb_i(n1,H) && inb_i(n2,W))) continue;
          float acc = 0;
          for (int c=0; c<F; c++){
            float g = go[(((int64_t)r0*F+c)*H+r1)*W+r2];
            int64_t iv = (((int64_t)n0*F+c)*H+n1)*W+n2;
            float add = g*weight;
            #pragma omp atomic
            gv[iv] += add;
            acc += g*v[iv];
          }
          gweights[fo] += acc;
        }
      }else{
        float c0 = bounds_f(ref[0]+inds_f[fo*3],T);
        float c1 = ref[1]+inds_f[fo*3+1], c2 = ref[2]+inds_f[fo*3+2];
        int signH = inb_f(c1,H) ? 1 : -1; c1 = bounds_f(c1,H);
        int signW = inb_f(c2,W) ? 1 : -1; c2 = bounds_f(c2,W);
        float n0 = bounds_f(c0,T);
        for (int pi=0; pi<ps; pi++) for (int pj=0; pj<ps; pj++){
          int r1 = ref[1]+dilation*(pi+patch_offset), r2 = ref[2]+dilation*(pj+patch_offset);
          float n1 = c1+(float)(dilation*(pi+patch_offset));
          int sH = inb_f(n1,H) ? 1 : -1; if (reflect_bounds) n1 = bounds_f(n1,H);
          float n2 = c2+(float)(dilation*(pj+patch_offset));
          int sW = inb_f(n2,W) ? 1 : -1; if (reflect_bounds) n2 = bounds_f(n2,W);
          if (!(inb_i(r0,T) && inb_i(r1,H) && inb_i(r2,W))) continue;
          if (!(inb_f(n0,T) && inb_f(n1,H) && inb_f(n2,W))) continue;
          int ti = (int)lrintf(roundf(n0));
          float acc=0, iW=0, iH=0;
          for (int c=0; c<F; c++){
            float g = go[(((int64_t)r0*F+c)*H+r1)*W+r2];
            float wg = g*weight, pix=0, igW=0, igH=0;
            const float* plane = v + ((int64_t)ti*F+c)*H*W;
            float* gplane = gv + ((int64_t)ti*F+c)*H*W;
            for (int ix=0; ix<2; ix++) for (int jx=0; jx<2; jx++){
              int h_i = (int)truncf(n1+ix), w_i = (int)truncf(n2+jx);
              float gH = fmaxf(0.f,1-fabsf(n1-h_i)), gW = fmaxf(0.f,1-fabsf(n2-w_i));
              float w = gH*gW;
              h_i = bounds_i(h_i,H); w_i = bounds_i(w_i,W);
              float vp = plane[(int64_t)h_i*W+w_i];
              pix += w*vp;
              igW += (jx==0) ? -gH*vp : gH*vp;
              igH += (ix==0) ? -gW*vp : gW*vp;
              float add = w*wg;
              #pragma omp atomic
              gplane[(int64_t)h_i*W+w_i] += add;
            }
            acc += g*pix;
            iW += g*weight*sW*igW;
            iH += g*weight*sH*igH;
          }
          gweights[fo] += acc;
          ginds[fo*3+1] += iH*signH;
          ginds[fo*3+2] += iW*signW;
        }
      }
    }
  }
}

/*
 * search_flow: chain frame-to-frame optical flows through the temporal window.
 *   forward : lib/csrc/nn/search_flow_kernel.cu:17-88 + lib/csrc/nn/shared_flows.cu:39-72
 *   backward: lib/csrc/nn/search_flow_kernel.cu:131-250 + lib/csrc/nn/shared_flows.cu:108-290
 * fflow,bflow (B,T,2,H,W); flows (B,T,W_t-1,2,nH,nW), W_t = min(2wt+1,T).
 * A tap still outside the frame after one reflection (an out-of-bounds read in the reference) is clamped.
 */
static inline int clamp_i(int v, int lim){ return v < 0 ? 0 : (v > lim-1 ? lim-1 : v); }

static void flow_hop(float* hc, float* wc, const float* flow, int H, int W){
  float h0 = *hc, w0 = *wc;
  for (int i=0;i<2;i++) for (int j=0;j<2;j++){
    int hj = (int)floorf(h0+i), wj = (int)floorf(w0+j);
    float weight = fmaxf(0.f,1-fabsf(hj-h0)) * fmaxf(0.f,1-fabsf(wj-w0));
    hj = clamp_i(bounds_i(hj,H),H); wj = clamp_i(bounds_i(wj,W),W);
    *wc = fmaf(weight, flow[(int64_t)hj*W+wj], *wc);
    *hc = fmaf(weight, flow[(int64_t)H*W+(int64_t)hj*W+wj], *hc);
  }
}

void oracle_search_flow_forward(float* flows, const float* fflow, const float* bflow,
                                int B, int T, int H, int W, int wt, int stride0){
  int nH=(H-1)/stride0+1, nW=(W-1)/stride0+1, nHW=nH*nW;
  int W_t = (2*wt+1 < T) ? 2*wt+1 : T;
  int64_t fs = 2LL*H*W;
  #pragma omp parallel for collapse(2) schedule(static)
  for (int b=0;b<B;b++) for (int qi=0; qi<T*nHW; qi++){
    int ref[3]; pixel_loc(ref,qi,stride0,nW,nHW,H,W);
    int ti=ref[0], hn=ref[1]/stride0, wn=ref[2]/stride0;
    float h_ref=(float)ref[1], w_ref=(float)ref[2], hc=h_ref, wc=w_ref;
    int t_max = time_tmax(ti,T,wt);
    for (int si=1; si<W_t; si++){
      int tj = ti+si;
      int swap = (ti+si-1)==t_max;
      tj = (tj>t_max) ? t_max-si : tj;
      if (swap){ hc=h_ref; wc=w_ref; }
      const float* flow = (tj>ti) ? fflow+((int64_t)b*T+(tj-1))*fs : bflow+((int64_t)b*T+(tj+1))*fs;
      flow_hop(&hc,&wc,flow,H,W);
      int64_t o = ((((int64_t)b*T+ti)*(W_t-1)+(si-1))*2)*nHW + (int64_t)hn*nW+wn;
      flows[o+nHW] = hc-h_ref;
      flows[o] = wc-w_ref;
    }
  }
}

static void flow_update_weights(float* dAdF0, float* dAdF1, float p1, float p2, int H, int W, const float* flow){
  float dF[2][2] = {{0,0},{0,0}};
  int sH = inb_f(p1,H)?1:-1, sW = inb_f(p2,W)?1:-1;
  p1 = bounds_f(p1,H); p2 = bounds_f(p2,W);
  for (int ix=0;ix<2;ix++) for (int jx=0;jx<2;jx++){
    int i0 = (int)(ix==0 ? floorf(p1) : floorf(p1)+1);
    float gH = fmaxf(0.f,1-fabsf(i0-p1));
    int i1 = (int)(jx==0 ? floorf(p2) : floorf(p2)+1);
    float gW = fmaxf(0.f,1-fabsf(i1-p2));
    if (!inb_i(i0,H) || !inb_i(i1,W)) continue;
    float vW = flow[(int64_t)i0*W+i1], vH = flow[(int64_t)H*W+(int64_t)i0*W+i1];
    dF[0][0] += (jx==0) ? -gH*vW : gH*vW;
    dF[0][1] += (ix==0) ? -gW*vW : gW*vW;
    dF[1][0] += (jx==0) ? -gH*vH : gH*vH;
    dF[1][1] += (ix==0) ? -gW*vH : gW*vH;
  }
  float a0=dAdF0[0], a1=dAdF0[1], b0=dAdF1[0], b1=dAdF1[1];
  dAdF0[0] += dF[0][0]*sW*a0 + dF[0][1]*sH*a1;
  dAdF0[1] += dF[1][0]*sW*a0 + dF[1][1]*sH*a1;
  dAdF1[0] += dF[0][0]*sW*b0 + dF[0][1]*sH*b1;
  dAdF1[1] += dF[1][0]*sW*b0 + dF[1][1]*sH*b1;
}

void oracle_search_flow_backward(float* gfflow, float* gbflow, const float* gflows,
                                 const float* fflow, const float* bflow, const float* flows,
                                 int B, int T, int H, int W, int wt, int stride0){
  int nH=(H-1)/stride0+1, nW=(W-1)/stride0+1, nHW=nH*nW;
  int W_t = (2*wt+1 < T) ? 2*wt+1 : T;
  int64_t fs = 2LL*H*W;
  /* serial over queries: the scatter targets overlap (the reference uses atomics) */
  for (int b=0;b<B;b++) for (int qi=0; qi<T*nHW; qi++) for (int si=1; si<=W_t; si++){
    int ref[3]; pixel_loc(ref,qi,stride0,nW,nHW,H,W);
    int ti=ref[0], nh=ref[1]/stride0, nw=ref[2]/stride0;
    int t_shift = (ti-wt<0 ? ti-wt : 0) + (ti+wt-(T-1)>0 ? ti+wt-(T-1) : 0);
    int t_min = ti-wt-t_shift; if (t_min<0) t_min=0;
    int t_max = ti+wt-t_shift; if (t_max>T-1) t_max=T-1;
    int tj = ti+si; tj = (tj>t_max) ? t_max-si : tj;
    int isFwd = tj>ti;
    int t_flow = isFwd ? tj-1 : tj+1;
    const float* flow = (isFwd?fflow:bflow)+(int64_t)b*T*fs;
    float* g_flow = (isFwd?gfflow:gbflow)+(int64_t)b*T*fs;
    int t_inc = isFwd?1:-1;
    int t_end = isFwd ? (t_max-tj+1) : (tj-t_min+1);
    if (t_end <= 0) continue;                    /* si == W_t: the reference thread does nothing */
    int adjacent = abs(ti-tj)==1;
    const float* fl = flows + (((int64_t)b*T+ti)*(W_t-1))*2*nHW + (int64_t)nh*nW+nw;
    const float* gf = gflows + (((int64_t)b*T+ti)*(W_t-1))*2*nHW + (int64_t)nh*nW+nw;
    float r1=(float)ref[1], r2=(float)ref[2];
    if (!adjacent){ r1 = ref[1]+fl[((int64_t)(si-2)*2+1)*nHW]; r2 = ref[2]+fl[((int64_t)(si-2)*2)*nHW]; }
    float dAdF0[2], dAdF1[2];
    for (int tx=0; tx<t_end; tx++){
      float gvW = gf[((int64_t)(si-1+tx)*2)*nHW], gvH = gf[((int64_t)(si-1+tx)*2+1)*nHW];
      if (tx==0){ dAdF0[0]=1; dAdF0[1]=0; dAdF1[0]=0; dAdF1[1]=1; }
      else{
        float p1 = (float)ref[1]+fl[((int64_t)(si+tx-2)*2+1)*nHW], p2 = (float)ref[2]+fl[((int64_t)(si+tx-2)*2)*nHW];
        flow_update_weights(dAdF0,dAdF1,p1,p2,H,W,flow+(int64_t)(t_flow+t_inc*tx)*fs);
      }
      float* g = g_flow+(int64_t)t_flow*fs;
      float v0 = dAdF0[0]*gvW+dAdF0[1]*gvH, v1 = dAdF1[0]*gvW+dAdF1[1]*gvH;
      for (int ix=0;ix<2;ix++) for (int jx=0;jx<2;jx++){     /* assign_bilin2d, shared_flows.cu:108-141 */
        int i0 = (int)floorf(r1+ix); float gH = fmaxf(0.f,1-fabsf(i0-r1));
        int i1 = (int)floorf(r2+jx); float gW = fmaxf(0.f,1-fabsf(i1-r2));
        float w = gH*gW;
        i0 = clamp_i(bounds_i(i0,H),H); i1 = clamp_i(bounds_i(i1,W),W);
        g[(int64_t)i0*W+i1] += w*v0;
        g[(int64_t)H*W+(int64_t)i0*W+i1] += w*v1;
      }
    }
  }
}
