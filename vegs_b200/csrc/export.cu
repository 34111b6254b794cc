// Parity support (not on the hot path): re-express this library's binning / geometry state in the
// reference's formats so tests can compare bit for bit with the reference's decoded
// geomBuffer / binningBuffer / imgBuffer.
//   point_list_keys[i] = (tile << 32) | depth bits     (rasterizer_impl.cu:70-111)
//   ranges[tile]       = [first, last)                  (rasterizer_impl.cu:116-138)
#include "common.cuh"

namespace vegs {

namespace {

__global__ void export_gaussians_kernel(int P, const int* __restrict__ radii,
                                        const float4* __restrict__ recA,
                                        const float4* __restrict__ recB,
                                        const float4* __restrict__ recC,
                                        const uint2* __restrict__ rect,
                                        uint32_t* __restrict__ tiles_touched,
                                        float* __restrict__ depths, float* __restrict__ means2D,
                                        float* __restrict__ conic_opacity, float* __restrict__ rgb) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const bool vis = radii[i] > 0;
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a, c = a;
    uint2 r = make_uint2(0u, 0u);
    if (vis) {
        const size_t gi = (size_t)kGeomRecStride * i;
        a = recA[gi]; b = recB[gi]; c = recC[gi]; r = rect[(size_t)kRectStride * i];
    }
    if (tiles_touched)
        tiles_touched[i] = ((r.x >> 16) - (r.x & 0xffffu)) * ((r.y >> 16) - (r.y & 0xffffu));
    if (depths) depths[i] = a.w;
    if (means2D) { means2D[2 * (size_t)i] = a.x; means2D[2 * (size_t)i + 1] = a.y; }
    if (conic_opacity) {
        float* o = conic_opacity + 4 * (size_t)i;
        o[0] = b.x; o[1] = b.y; o[2] = b.z; o[3] = b.w;
    }
    if (rgb) { float* o = rgb + 3 * (size_t)i; o[0] = c.x; o[1] = c.y; o[2] = c.z; }
}

__global__ void export_keys_kernel(const uint2* __restrict__ ranges,
                                   const uint32_t* __restrict__ point_list,
                                   const float4* __restrict__ recA, uint32_t id_mask,
                                   uint64_t* __restrict__ keys, uint32_t* __restrict__ list_out) {
    const uint32_t tile = blockIdx.x;
    const uint2 rg = ranges[tile];
    for (uint32_t i = rg.x + threadIdx.x; i < rg.y; i += blockDim.x) {
        const uint32_t g = point_list[i] & id_mask;   // strip the unit mask byte (common.cuh)
        if (keys) keys[i] = ((uint64_t)tile << 32) | (uint64_t)__float_as_uint(recA[(size_t)kGeomRecStride * g].w);
        if (list_out) list_out[i] = g;
    }
}

}  // namespace

int launch_export(const VegsForwardArgs& a, const GeomView& g, const ScratchView& s,
                  const ImgView& im, const BinView& b, uint32_t R, uint64_t* keys,
                  uint32_t* point_list, uint32_t* ranges, uint32_t* tiles_touched, float* depths,
                  float* means2D, float* conic_opacity, float* rgb, uint32_t* n_contrib,
                  cudaStream_t stream) {
    const int T = ((a.W + kTile - 1) / kTile) * ((a.H + kTile - 1) / kTile);
    if (a.P > 0 && (tiles_touched || depths || means2D || conic_opacity || rgb)) {
        export_gaussians_kernel<<<(a.P + 255) / 256, 256, 0, stream>>>(
            a.P, a.radii, g.recA, g.recB, g.recC, s.rect, tiles_touched, depths, means2D,
            conic_opacity, rgb);
        VEGS_LAUNCH_CHECK("export_gaussians", stream, true);
    }
    if (R > 0 && (keys || point_list)) {
        export_keys_kernel<<<T, 256, 0, stream>>>(im.ranges, b.point_list, g.recA,
                                                  list_has_unit_masks(a.P) ? kListIdMask : 0xffffffffu, keys,
                                                  point_list);
        VEGS_LAUNCH_CHECK("export_keys", stream, true);
    }
    if (ranges)
        VEGS_CUDA_OK(cudaMemcpyAsync(ranges, im.ranges, (size_t)T * sizeof(uint2),
                                     cudaMemcpyDeviceToDevice, stream));
    if (n_contrib)
        VEGS_CUDA_OK(cudaMemcpyAsync(n_contrib, im.n_contrib, (size_t)a.W * a.H * sizeof(uint32_t),
                                     cudaMemcpyDeviceToDevice, stream));
    return VEGS_OK;
}

}  // namespace vegs
