// extern "C" entry points of libvegs_b200.so (declared in include/vegs_b200.h) and the host-side
// launch sequence.  Replaces CudaRasterizer::Rasterizer::forward / backward / markVisible
// (reference cuda_rasterizer/rasterizer_impl.cu:141-153, :197-350, :356-481) and the torch binding
// layer above them (rasterize_points.cu).  All buffers are caller-provided; nothing here
// allocates device memory.
#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <vector>
#include "common.cuh"

namespace vegs {

static thread_local char g_error[512] = "";
static std::atomic<uint64_t> g_launches{0};

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_error, sizeof(g_error), fmt, ap);
    va_end(ap);
}

void count_launch(int n) { g_launches.fetch_add((uint64_t)n, std::memory_order_relaxed); }

// ---- optional per-kernel timing with CUDA events on the launching stream (bench.py's roofline
// leg).  When enabled, an event is recorded after every launch; a kernel's time is the interval
// from the previous event of the same entry-point call to its own.
struct ProfEvent { const char* name; cudaEvent_t ev; };
static std::atomic<bool> g_prof_on{false};
static std::mutex g_prof_mu;
static std::vector<ProfEvent> g_prof_events;

void prof_mark(const char* name, cudaStream_t stream) {
    if (!g_prof_on.load(std::memory_order_relaxed)) return;
    cudaEvent_t ev;
    if (cudaEventCreate(&ev) != cudaSuccess) return;
    cudaEventRecord(ev, stream);
    std::lock_guard<std::mutex> lk(g_prof_mu);
    g_prof_events.push_back({name, ev});
}

// 32-bit depth keys, kRadixBits per pass: the sorted order ends in the second buffer pair iff the
// number of passes is odd.  Used by both forward stages so they cannot disagree.
static constexpr bool depth_sort_in_second() { return (((32 + kRadixBits - 1) / kRadixBits) & 1) != 0; }

static int tile_bits(int T) {
    int b = 0;
    while ((1 << b) < T) b++;
    return b;  // ids 0..T-1 fit in b bits
}

static int validate_common(int P, int D, int M, int W, int H) {
    if (P < 0 || W <= 0 || H <= 0 || D < 0 || D > 3) {
        set_error("bad sizes: P=%d W=%d H=%d D=%d", P, W, H, D);
        return VEGS_EINVAL;
    }
    if (M != 0 && M < (D + 1) * (D + 1)) {
        set_error("M=%d SH coefficients cannot hold degree %d", M, D);
        return VEGS_EINVAL;
    }
    if (W > 65535 || H > 65535) {
        set_error("image too large: tile rectangles and pixel boxes are packed in 16-bit coordinates (W, H <= 65535)");
        return VEGS_EINVAL;
    }
    return VEGS_OK;
}

}  // namespace vegs

using namespace vegs;

extern "C" int vegs_abi_version(void) { return VEGS_ABI_VERSION; }
extern "C" const char* vegs_last_error(void) { return g_error; }
extern "C" uint64_t vegs_kernel_launch_count(void) { return g_launches.load(); }

extern "C" void vegs_profile_enable(int32_t on) { g_prof_on.store(on != 0); }

static std::atomic<bool> g_unit_masks{true};
extern "C" void vegs_debug_unit_masks(int32_t on) { g_unit_masks.store(on != 0); }
namespace vegs {
bool sh_bulk_disabled() {
    static const bool off = [] { const char* e = getenv("VEGS_SH_BULK"); return e != nullptr && e[0] == '0'; }();
    return off;
}
bool list_has_unit_masks(int P) { return g_unit_masks.load() && P <= (int)kListIdMask; }
}

extern "C" int vegs_profile_collect(char* buf, size_t buf_bytes) {
    std::vector<ProfEvent> evs;
    {
        std::lock_guard<std::mutex> lk(g_prof_mu);
        evs.swap(g_prof_events);
    }
    std::map<std::string, std::pair<double, long>> acc;
    cudaError_t err = cudaSuccess;
    for (size_t i = 0; i < evs.size(); i++) {
        if (err == cudaSuccess) err = cudaEventSynchronize(evs[i].ev);
        if (err == cudaSuccess && i > 0 && std::string(evs[i].name) != "__start") {
            float ms = 0.f;
            if (cudaEventElapsedTime(&ms, evs[i - 1].ev, evs[i].ev) == cudaSuccess) {
                auto& e = acc[evs[i].name];
                e.first += ms;
                e.second += 1;
            }
        }
    }
    for (auto& e : evs) cudaEventDestroy(e.ev);
    if (err != cudaSuccess) { set_error("profile collect: %s", cudaGetErrorString(err)); return VEGS_ECUDA; }
    std::string out;
    char line[256];
    for (auto& kv : acc) {
        snprintf(line, sizeof(line), "%s %.6f %ld\n", kv.first.c_str(), kv.second.first, kv.second.second);
        out += line;
    }
    if (!buf || buf_bytes < out.size() + 1) { set_error("profile buffer too small"); return VEGS_ESIZE; }
    memcpy(buf, out.c_str(), out.size() + 1);
    return VEGS_OK;
}

extern "C" size_t vegs_geom_bytes(int32_t P) { return geom_view(nullptr, P).bytes + 256; }
extern "C" size_t vegs_scratch_bytes(int32_t P) { return scratch_view(nullptr, P).bytes + 256; }
extern "C" size_t vegs_img_bytes(int32_t W, int32_t H) { return img_view(nullptr, W, H).bytes + 256; }
extern "C" size_t vegs_binning_bytes(int64_t R) { return bin_view(nullptr, R).bytes + 256; }
extern "C" size_t vegs_bin_scratch_bytes(int64_t R, int32_t W, int32_t H) {
    (void)W; (void)H;
    return bin_scratch_view(nullptr, R).bytes + 256;
}
extern "C" size_t vegs_bwd_scratch_bytes(int32_t P) {
    // 48 B per Gaussian of blend-stage accumulators + 32 B for the opt-in live map gradients
    return align_up((size_t)(P > 0 ? P : 1) * 48) + align_up((size_t)(P > 0 ? P : 1) * 32) + 256;
}

static inline void* aligned(const void* p) {
    return reinterpret_cast<void*>(align_up(reinterpret_cast<uintptr_t>(p)));
}

extern "C" int vegs_forward_prepare(const VegsForwardArgs* a, void* stream_) {
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    if (!a) { set_error("null args"); return VEGS_EINVAL; }
    int rc = validate_common(a->P, a->D, a->M, a->W, a->H);
    if (rc) return rc;
    if (!a->num_rendered_host) { set_error("num_rendered_host is null"); return VEGS_EINVAL; }
    if (a->P == 0) { a->num_rendered_host[0] = 0; a->num_rendered_host[1] = 0; return VEGS_OK; }
    if (!a->means3D || !a->opacities || !a->viewmatrix || !a->projmatrix || !a->campos ||
        !a->radii || !a->geom || !a->scratch) {
        set_error("vegs_forward_prepare: a required pointer is null");
        return VEGS_EINVAL;
    }
    if ((a->shs == nullptr) == (a->colors_precomp == nullptr)) {
        set_error("provide exactly one of shs / colors_precomp");
        return VEGS_EINVAL;
    }
    const bool have_sr = a->scales != nullptr && a->rotations != nullptr;
    if (have_sr == (a->cov3D_precomp != nullptr) ||
        ((a->scales != nullptr) != (a->rotations != nullptr))) {
        set_error("provide exactly one of (scales, rotations) / cov3D_precomp");
        return VEGS_EINVAL;
    }
    if (a->shs && a->M <= 0) { set_error("shs given but M == 0"); return VEGS_EINVAL; }

    const GeomView g = geom_view(aligned(a->geom), a->P);
    const ScratchView s = scratch_view(aligned(a->scratch), a->P);

    prof_mark("__start", stream);
    rc = launch_preprocess_fwd(*a, g, s, stream);
    if (rc) return rc;
    bool second = false;
    rc = radix_sort_pairs(s.keys0, s.vals0, s.keys1, s.vals1, /*implicit_vals=*/true,
                          (uint32_t)a->P, 32, s.hist, s.rowtot, /*write_last_keys=*/false, &second,
                          stream, a->debug != 0);
    if (rc) return rc;
    if (second != depth_sort_in_second()) { set_error("internal: depth sort pass parity"); return VEGS_EINVAL; }
    const uint32_t* order = second ? s.vals1 : s.vals0;
    rc = launch_tile_count(a->P, order, s.rect, s.chunksum, s.partbase, g.counters, stream,
                           a->debug != 0);
    if (rc) return rc;
    VEGS_CUDA_OK(cudaMemcpyAsync(a->num_rendered_host, g.counters, 2 * sizeof(int32_t),
                                 cudaMemcpyDeviceToHost, stream));
    prof_mark("num_rendered_d2h", stream);
    return VEGS_OK;
}

extern "C" int vegs_forward_render(const VegsForwardArgs* a, int32_t num_rendered, void* stream_) {
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    if (!a) { set_error("null args"); return VEGS_EINVAL; }
    int rc = validate_common(a->P, a->D, a->M, a->W, a->H);
    if (rc) return rc;
    if (a->P == 0) return VEGS_OK;
    if (num_rendered < 0) { set_error("num_rendered < 0"); return VEGS_EINVAL; }
    if (!a->background || !a->out_color || !a->out_depth || !a->out_cov_quat ||
        !a->out_cov_scale || !a->out_alpha || !a->geom || !a->scratch || !a->img ||
        (num_rendered > 0 && (!a->binning || !a->bin_scratch))) {
        set_error("vegs_forward_render: a required pointer is null");
        return VEGS_EINVAL;
    }
    const bool debug = a->debug != 0;
    const uint32_t R = (uint32_t)num_rendered;
    const int gx = (a->W + kTile - 1) / kTile, gy = (a->H + kTile - 1) / kTile;
    const int T = gx * gy;
    const GeomView g = geom_view(aligned(a->geom), a->P);
    const ScratchView s = scratch_view(aligned(a->scratch), a->P);
    const ImgView im = img_view(aligned(a->img), a->W, a->H);
    BinView b;
    b.point_list = nullptr;
    b.bytes = 0;
    const uint32_t* sorted_keys = nullptr;
    prof_mark("__start", stream);
    if (R > 0) {
        b = bin_view(aligned(a->binning), R);
        const BinScratchView bs = bin_scratch_view(aligned(a->bin_scratch), R);
        const int nbits = tile_bits(T);
        const int passes = (nbits + kRadixBits - 1) / kRadixBits;
        // Arrange the ping-pong so the sorted values land in the persistent point_list.
        uint32_t *k0 = bs.tkeys0, *k1 = bs.tkeys1, *v0, *v1;
        if ((passes & 1) == 0) { v0 = b.point_list; v1 = bs.tvals1; }
        else { v0 = bs.tvals1; v1 = b.point_list; }
        // the depth sort's result buffer, derived exactly as vegs_forward_prepare derives it
        const uint32_t* order = depth_sort_in_second() ? s.vals1 : s.vals0;
        // Upper bound of the emission work items (chunk c needs ceil(instances_c / part) CTAs): the
        // exact count lives on the device (partbase[nchunks]); surplus CTAs exit at once, so the
        // host never has to read it back.
        const uint32_t nparts = div_up((uint32_t)a->P, kEmitChunk) + div_up(R, kEmitPartInstances);
        rc = launch_emit(a->P, order, s.rect, s.chunksum, s.partbase, nparts, gx, k0, v0, R, s.pbox, a->W, a->H,
                         stream, debug);
        if (rc) return rc;
        bool second = false;
        rc = radix_sort_pairs(k0, v0, k1, v1, false, R, nbits, bs.hist, bs.rowtot,
                              /*write_last_keys=*/true, &second, stream, debug);
        if (rc) return rc;
        sorted_keys = second ? k1 : k0;
    }
    rc = launch_tile_ranges(R, sorted_keys, im.ranges, im.tile_order, T, stream, debug);
    if (rc) return rc;
    return launch_composite_fwd(*a, g, s, im, b, stream);
}

extern "C" int vegs_backward(const VegsBackwardArgs* a, void* stream_) {
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    if (!a) { set_error("null args"); return VEGS_EINVAL; }
    int rc = validate_common(a->P, a->D, a->M, a->W, a->H);
    if (rc) return rc;
    if (a->P == 0) return VEGS_OK;
    if (!a->means3D || !a->viewmatrix || !a->projmatrix || !a->campos || !a->radii ||
        !a->out_alpha || !a->background || !a->geom || !a->img || !a->bwd_scratch ||
        (a->R > 0 && !a->binning)) {
        set_error("vegs_backward: a required pointer is null");
        return VEGS_EINVAL;
    }
    if (a->shs && !a->dL_dsh && !a->dL_drgb_deferred) { set_error("shs given but neither dL_dsh nor dL_drgb_deferred"); return VEGS_EINVAL; }
    if ((a->scales != nullptr) != (a->rotations != nullptr)) {
        set_error("scales and rotations must be given together");
        return VEGS_EINVAL;
    }
    const GeomView g = geom_view(aligned(a->geom), a->P);
    const ImgView im = img_view(aligned(a->img), a->W, a->H);
    BinView b;
    b.point_list = nullptr;
    b.bytes = 0;
    if (a->R > 0) b = bin_view(aligned(a->binning), a->R);
    float4* acc = reinterpret_cast<float4*>(aligned(a->bwd_scratch));
    prof_mark("__start", stream);
    VEGS_CUDA_OK(cudaMemsetAsync(acc, 0, (size_t)a->P * 48, stream));
    prof_mark("memset_bwd_acc", stream);
    float4* map_acc = nullptr;
    if (a->live_map_gradients) {
        if (!a->out_cov_quat || a->cov3D_precomp) {
            set_error("vegs_backward: live_map_gradients needs out_cov_quat and the scale / rotation inputs");
            return VEGS_EINVAL;
        }
        map_acc = reinterpret_cast<float4*>(reinterpret_cast<char*>(acc) + align_up((size_t)a->P * 48));
        VEGS_CUDA_OK(cudaMemsetAsync(map_acc, 0, (size_t)a->P * 32, stream));
    }
    if (a->R > 0) {
        rc = launch_composite_bwd(*a, g, im, b, acc, map_acc, stream);
        if (rc) return rc;
    }
    return launch_preprocess_bwd(*a, g, acc, map_acc, stream);
}

extern "C" int vegs_mark_visible(int32_t P, const float* means3D, const float* viewmatrix,
                                 const float* projmatrix, uint8_t* present, void* stream_) {
    (void)projmatrix;  // the reference computes an unused projection (auxiliary.h:148-152)
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    if (P < 0 || (P > 0 && (!means3D || !viewmatrix || !present))) {
        set_error("vegs_mark_visible: bad argument");
        return VEGS_EINVAL;
    }
    if (P == 0) return VEGS_OK;
    return launch_mark_visible(P, means3D, viewmatrix, present, stream);
}

extern "C" int vegs_export_state(const VegsForwardArgs* a, int32_t num_rendered,
                                 uint64_t* point_list_keys, uint32_t* point_list, uint32_t* ranges,
                                 uint32_t* tiles_touched, float* depths, float* means2D,
                                 float* conic_opacity, float* rgb, uint32_t* n_contrib,
                                 void* stream_) {
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    if (!a || a->P <= 0 || !a->geom || !a->scratch || !a->img || !a->radii) {
        set_error("vegs_export_state: bad argument");
        return VEGS_EINVAL;
    }
    const GeomView g = geom_view(aligned(a->geom), a->P);
    const ScratchView s = scratch_view(aligned(a->scratch), a->P);
    const ImgView im = img_view(aligned(a->img), a->W, a->H);
    BinView b;
    b.point_list = nullptr;
    b.bytes = 0;
    if (num_rendered > 0) {
        if (!a->binning) { set_error("vegs_export_state: binning is null"); return VEGS_EINVAL; }
        b = bin_view(aligned(a->binning), num_rendered);
    }
    return launch_export(*a, g, s, im, b, (uint32_t)num_rendered, point_list_keys, point_list,
                         ranges, tiles_touched, depths, means2D, conic_opacity, rgb, n_contrib,
                         stream);
}

extern "C" int vegs_export_tile_clocks(const void* img, int32_t W, int32_t H, uint32_t* fwd, uint32_t* bwd,
                                       uint32_t* dbg, void* stream_) {
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    if (!img || W <= 0 || H <= 0) { set_error("vegs_export_tile_clocks: bad argument"); return VEGS_EINVAL; }
    const ImgView im = img_view(aligned(img), W, H);
    const size_t T = (size_t)((W + kTile - 1) / kTile) * ((H + kTile - 1) / kTile);
    if (fwd) VEGS_CUDA_OK(cudaMemcpyAsync(fwd, im.tile_clocks, T * 4, cudaMemcpyDeviceToDevice, stream));
    if (bwd) VEGS_CUDA_OK(cudaMemcpyAsync(bwd, im.tile_clocks + T, T * 4, cudaMemcpyDeviceToDevice, stream));
    if (dbg) VEGS_CUDA_OK(cudaMemcpyAsync(dbg, im.tile_dbg, T * 64, cudaMemcpyDeviceToDevice, stream));
    return VEGS_OK;
}
