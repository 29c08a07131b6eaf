// compact.cu — the 32-byte binary node (NodeQ, kernels.h): the grid and the packing kernel.
//
// Why: the traversal is bound by the L1 load pipe (ncu: LSU wavefronts at 0.73-0.81 of peak,
// DRAM at 0.07), and a node step of the 64-byte Node costs two 32-byte requests per lane.
// Sixteen-bit planes on one grid per scene halve that, and unlike the per-node frames of the
// 8-wide layout (wide8.cu) the decode costs nothing: the grid is the same for every node, so
// origin, step and the ray fold into per-RAY constants and a plane distance is one PRMT (the
// 16 bits become the mantissa of 2^23 + q) and one FFMA.
//
// Exactness: the traversal only has to be CONSERVATIVE (trace_common.cuh, acceptance rule): it
// may never cull a node that holds a triangle whose own box passes slab().  Packing rounds every
// plane outwards by at least one whole step, which covers the rounding of the per-ray constants
// ("compact nodes"); the hits are those of the 64-byte layout bit for bit.
#include <math.h>

#include "common.cuh"
#include "kernels.h"

namespace rt {

// Grid of one scene: 65,000 steps across the box, three more below it; the step is never finer
// than 4 ulp of the box's largest coordinate, so that rounding `origin` moves the grid by less
// than an eighth of a step.
QGrid make_qgrid(const float b[6]) {
  QGrid g;
  float* o = &g.ox;
  float* st = &g.sx;
  for (int a = 0; a < 3; ++a) {
    const double lo = b[a], hi = b[a + 3];
    const double big = fmax(fabs(lo), fabs(hi));
    double step = (hi - lo) / 65000.0;
    step = fmax(step, big * 4.76837158203125e-07);  // 4 ulp = 2^-21 of the magnitude
    step = fmax(step, 1e-30);
    float sf = (float)step;
    if ((double)sf < step) sf = nextafterf(sf, INFINITY);
    st[a] = sf;
    o[a] = (float)(lo - 3.0 * (double)sf);
  }
  return g;
}

namespace {

// outward 16-bit plane pair of [lo, hi] on one axis: at least one whole step beyond the box
__device__ __forceinline__ uint32_t pack_axis(float lo, float hi, float origin, float step) {
  const double inv = 1.0 / (double)step;
  const double xl = ((double)lo - (double)origin) * inv, xh = ((double)hi - (double)origin) * inv;
  int ql = (int)floor(xl - 1e-6) - 1, qh = (int)ceil(xh + 1e-6) + 1;
  ql = max(0, min(65535, ql));
  qh = max(0, min(65535, qh));
  return (uint32_t)ql | ((uint32_t)qh << 16);
}

__global__ void __launch_bounds__(256)
pack_nodes_kernel(const Node* __restrict__ nodes, uint32_t n_nodes, QGrid g, NodeQ* __restrict__ out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_nodes) return;
  const float4* p = reinterpret_cast<const float4*>(nodes + i);
  const float4 lo0 = ld_stream(p), hi0 = ld_stream(p + 1), lo1 = ld_stream(p + 2), hi1 = ld_stream(p + 3);
  uint4 a, b;
  a.x = pack_axis(lo0.x, hi0.x, g.ox, g.sx);
  a.y = pack_axis(lo0.y, hi0.y, g.oy, g.sy);
  a.z = pack_axis(lo0.z, hi0.z, g.oz, g.sz);
  a.w = pack_axis(lo1.x, hi1.x, g.ox, g.sx);
  b.x = pack_axis(lo1.y, hi1.y, g.oy, g.sy);
  b.y = pack_axis(lo1.z, hi1.z, g.oz, g.sz);
  b.z = __float_as_uint(lo0.w);
  b.w = __float_as_uint(lo1.w);
  uint4* q = reinterpret_cast<uint4*>(out + i);
  q[0] = a;
  q[1] = b;
}

}  // namespace

int launch_pack_nodes(const Node* nodes, uint64_t n_nodes, QGrid grid, NodeQ* out, cudaStream_t s) {
  if (n_nodes == 0) return 0;
  pack_nodes_kernel<<<(unsigned)((n_nodes + 255) / 256), 256, 0, s>>>(nodes, (uint32_t)n_nodes, grid, out);
  RT_LAUNCHED();
  return 0;
}

}  // namespace rt
