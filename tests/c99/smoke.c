/* C99 translation unit that includes the public header and drives the three entry
 * points (build_bvh / trace_rays / free_bvh): the boundary compiles and links from plain C.
 *   no CUDA device: build_bvh must fail with RT_ERR_NO_DEVICE (exit 0, prints "no device")
 *   with a device:  one triangle, two rays: one hit at t = 1, one miss (exit 0, "ok") */
#include <math.h>
#include <stdio.h>
#include <string.h>

#include "gpu_raytracer.h"

int main(void) {
  const float tris[12] = {0, 0, 1, 0, 1, 0, 1, 0, 0, 1, 1, 0};
  const float rays[16] = {0.25f, 0.25f, 0, 0, 0, 0, 1, INFINITY, 5, 5, 0, 0, 0, 0, 1, INFINITY};
  float hits[8];
  rt_build_opts opts;
  rt_bvh* h = NULL;
  rt_bvh_info info;
  int prim0, prim1, rc;
  uint64_t first, count;

  memset(&opts, 0, sizeof(opts));
  opts.key_bits = 30;
  opts.device = -1;
  if (rt_shard_range(100, 1, 3, &first, &count) != RT_OK || first != 64 || count != 32) {
    printf("rt_shard_range: %llu %llu\n", (unsigned long long)first, (unsigned long long)count);
    return 1;
  }
  rc = build_bvh(tris, 1, &opts, &h);
  if (rt_device_count() == 0) {
    if (rc != RT_ERR_NO_DEVICE || h != NULL) {
      printf("expected RT_ERR_NO_DEVICE, got %d\n", rc);
      return 1;
    }
    printf("no device: %s\n", bvh_last_error());
    return 0;
  }
  if (rc != RT_OK) {
    printf("build_bvh: %d %s\n", rc, bvh_last_error());
    return 1;
  }
  if (bvh_info(h, &info) != RT_OK || info.n_tris != 1) return 1;
  rc = trace_rays(h, rays, 2, RT_TRACE_CLOSEST, hits);
  if (rc != RT_OK) {
    printf("trace_rays: %d %s\n", rc, bvh_last_error());
    return 1;
  }
  memcpy(&prim0, &hits[3], 4);
  memcpy(&prim1, &hits[7], 4);
  free_bvh(h);
  if (prim0 != 0 || hits[0] != 1.0f || prim1 != -1) {
    printf("wrong hits: prim %d t %g, prim %d\n", prim0, hits[0], prim1);
    return 1;
  }
  printf("ok\n");
  return 0;
}
