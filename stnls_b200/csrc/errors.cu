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
