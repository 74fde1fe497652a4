// error text + small host helpers of the C ABI
#include <cstdarg>
#include <cstdio>
#include "common.cuh"

static thread_local char g_err[512] = "";

void stnls_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

extern "C" const char* stnls_last_error(void) { return g_err; }
extern "C" int stnls_version(void) { return 100; }

// separable patch-overlap counts; equals the atomically counted map of
// lib/csrc/agg/gather_add_kernel.cu:88-97 (ref pixel = query pixel + dilation*(p+patch_offset), never reflected)
extern "C" int stnls_overlap_counts(int* counts, int H, int nH, int ps, int stride0, int dilation, int patch_offset) {
  for (int h = 0; h < H; h++) counts[h] = 0;
  for (int nh = 0; nh < nH; nh++) {
    int ref = (nh * stride0) % H;
    for (int pi = 0; pi < ps; pi++) {
      int h = ref + dilation * (pi + patch_offset);
      if (h >= 0 && h <= H - 1) counts[h] += 1;
    }
  }
  int mn = 1 << 30;
  for (int h = 0; h < H; h++) mn = counts[h] < mn ? counts[h] : mn;
  return mn;
}

// frame-shard consistency (see common.cuh).  Window rule: lib/csrc/shared_kernel.cu:300-312.
int stnls_check_geom(const stnls_geom* g, int wt, const char* who) {
  STNLS_REQUIRE(g->B >= 0 && g->HD > 0 && g->T >= 0 && g->Tq >= 0 && g->Tk >= 0, "%s: negative sizes", who);
  STNLS_REQUIRE(g->tq0 >= 0 && g->tq0 + g->Tq <= g->T, "%s: query frames [%d,%d) outside the clip [0,%d)", who, g->tq0,
                g->tq0 + g->Tq, g->T);
  STNLS_REQUIRE(g->tk0 >= 0 && g->tk0 + g->Tk <= g->T, "%s: key frames [%d,%d) outside the clip [0,%d)", who, g->tk0,
                g->tk0 + g->Tk, g->T);
  if (wt >= 0 && g->Tq > 0) {
    int lo = g->T, hi = -1;
    for (int t = g->tq0; t < g->tq0 + g->Tq; t++) {
      const int sh = (t - wt < 0 ? t - wt : 0) + (t + wt - (g->T - 1) > 0 ? t + wt - (g->T - 1) : 0);
      const int a = t - wt - sh > 0 ? t - wt - sh : 0;
      const int b = t + wt - sh < g->T - 1 ? t + wt - sh : g->T - 1;
      lo = a < lo ? a : lo;
      hi = b > hi ? b : hi;
    }
    STNLS_REQUIRE(lo >= g->tk0 && hi < g->tk0 + g->Tk,
                  "%s: resident key frames [%d,%d) do not cover the temporal windows [%d,%d] of query frames [%d,%d)", who,
                  g->tk0, g->tk0 + g->Tk, lo, hi, g->tq0, g->tq0 + g->Tq);
  }
  return STNLS_OK;
}
