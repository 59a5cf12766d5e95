// module.cpp -- `_vulkan_forge`: the Python-visible drop-in surface (pybind11) above the C ABI.
// Mirrors the PyO3 module of the reference (src/lib.rs:961-976): same class / function names,
// argument names, defaults, exception types and message strings.  All GPU work goes through
// include/vf_cuda.h (libvfcuda.so); there is no CPU fallback -- without a CUDA device the
// constructors raise RuntimeError("No suitable GPU adapter ...") like src/terrain/mod.rs:285.
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include "../../../include/vf_cuda.h"
#include "vf_camera.hpp"

namespace py = pybind11;
using vfcam::M4;
using vfcam::V3;

// ---- error mapping ---------------------------------------------------------------------
static void check(int rc) {
    if (rc == VF_OK) return;
    const std::string msg = vf_last_error();
    if (rc == VF_E_OOM) throw std::runtime_error("out of memory: " + msg);
    throw std::runtime_error(msg);
}

// ---- pinned-host pool: numpy results are built directly on pinned memory so the readback is a
// single DMA (the reference un-pads rows on the CPU, src/lib.rs:204-211) -------------------
namespace pool {
static std::mutex mu;
static std::multimap<size_t, void*> free_list;
static size_t cached_bytes = 0;
// pinned blocks kept for reuse (VF_PINNED_POOL_MB, default 4096): batches come back in blocks of hundreds of MB, and
// cudaHostAlloc / cudaFreeHost of such a block costs tens of milliseconds
static const size_t kMaxCached = [] { const char* e = std::getenv("VF_PINNED_POOL_MB"); return (size_t)(e ? std::atoll(e) : 4096) << 20; }();

static void* take(size_t bytes) {
    {
        std::lock_guard<std::mutex> lk(mu);
        auto it = free_list.find(bytes);
        if (it != free_list.end()) {
            void* p = it->second;
            free_list.erase(it);
            cached_bytes -= bytes;
            return p;
        }
    }
    void* p = nullptr;
    check(vf_host_alloc(bytes, &p));
    return p;
}
static void give(void* p, size_t bytes) {
    {
        std::lock_guard<std::mutex> lk(mu);
        if (cached_bytes + bytes <= kMaxCached) {
            free_list.emplace(bytes, p);
            cached_bytes += bytes;
            return;
        }
    }
    vf_host_free(p);
}
struct Block { void* p; size_t bytes; };
}  // namespace pool

template <typename T>
static py::array_t<T> pinned_array(const std::vector<py::ssize_t>& shape) {
    size_t n = 1;
    for (auto d : shape) n *= (size_t)d;
    const size_t bytes = std::max<size_t>(n * sizeof(T), 1);
    auto* blk = new pool::Block{pool::take(bytes), bytes};
    py::capsule owner(blk, [](void* q) {
        auto* b = static_cast<pool::Block*>(q);
        pool::give(b->p, b->bytes);
        delete b;
    });
    return py::array_t<T>(shape, static_cast<T*>(blk->p), owner);
}

static V3 v3(const std::tuple<float, float, float>& t) { return {std::get<0>(t), std::get<1>(t), std::get<2>(t)}; }

// src/camera.rs:94-112: column-major Mat4 -> (4,4) float32, M[row, col]
static py::array_t<float> mat4_to_numpy(const M4& m) {
    py::array_t<float> a({4, 4});
    auto r = a.mutable_unchecked<2>();
    for (int row = 0; row < 4; ++row)
        for (int col = 0; col < 4; ++col) r(row, col) = m.m[4 * col + row];
    return a;
}

// ---- colormap registry (src/colormap/mod.rs:7-34) ---------------------------------------
static const char* kSupported[3] = {"viridis", "magma", "terrain"};
static int colormap_id(const std::string& name) {
    for (int i = 0; i < 3; ++i)
        if (name == kSupported[i]) return i;
    throw std::runtime_error("Unknown colormap '" + name + "'. Supported: viridis, magma, terrain");
}

// ---- Globals / uniforms (src/terrain/mod.rs:114-215) ------------------------------------
struct Globals {
    V3 sun_dir = vfcam::normalize({0.5f, 0.8f, 0.6f});
    float exposure = 1.0f, spacing = 1.0f, h_min = -0.5f, h_max = 0.5f, exaggeration = 1.0f;
    void to_uniforms(const M4& view, const M4& proj, float out[44]) const {
        std::memcpy(out, view.m, 64);
        std::memcpy(out + 16, proj.m, 64);
        out[32] = sun_dir.x; out[33] = sun_dir.y; out[34] = sun_dir.z; out[35] = exposure;
        out[36] = spacing; out[37] = h_max - h_min; out[38] = exaggeration; out[39] = 0.0f;
        out[40] = out[41] = out[42] = out[43] = 0.0f;
    }
};

// src/lib.rs:444-453
static V3 sun_from_angles(float elevation_deg, float azimuth_deg) {
    const float pi = 3.14159265358979323846f;
    const float el = elevation_deg * pi / 180.0f, az = azimuth_deg * pi / 180.0f;
    const float se = std::sin(el), ce = std::cos(el), sa = std::sin(az), ca = std::cos(az);
    return vfcam::normalize_or_zero({ce * ca, se, ce * sa});
}

// ========================================================================================
// Terrain objects: Scene (src/scene/mod.rs) and TerrainSpike (src/terrain/mod.rs:221-547)
// share everything but their defaults.
class TerrainObject {
public:
    TerrainObject(uint32_t width, uint32_t height, py::object grid, py::object colormap, bool spike)
        : width_(width), height_(height), spike_(spike) {
        uint32_t g = grid.is_none() ? 128u : grid.cast<uint32_t>();
        grid_ = std::max(g, 2u);                                           // scene/mod.rs:61
        const std::string cmap = colormap.is_none() ? "viridis" : colormap.cast<std::string>();
        const int lut = colormap_id(cmap);                                 // RuntimeError if unknown
        // R2a runtime format selection, src/terrain/mod.rs:45-49 (sRGB sampling always available here)
        lut_unorm_ = std::getenv("VF_FORCE_LUT_UNORM") != nullptr;
        if (width == 0 || height == 0) throw std::runtime_error("width and height must be > 0");
        check(vf_scene_create((int)width, (int)height, grid_, lut, lut_unorm_ ? 1 : 0, &scene_));
        if (spike) globals_.sun_dir = vfcam::normalize({0.5f, 1.0f, 0.3f});  // terrain/mod.rs:325-327,689
        view_ = vfcam::look_at_rh({3.0f, 2.0f, 3.0f}, {0.0f, 0.0f, 0.0f}, {0.0f, 1.0f, 0.0f});
        proj_ = vfcam::perspective_wgpu(vfcam::to_radians(45.0f), (float)width / (float)height, 0.1f, 100.0f);
        push_uniforms();
        if (spike) {
            const float z = 0.0f;                                          // 1x1 dummy, terrain/mod.rs:342-378
            check(vf_scene_set_height_f32(scene_, &z, 1, 1));
        } else {
            const float d[4] = {0.00f, 0.25f, 0.50f, 0.75f};               // 2x2 dummy, scene/mod.rs:142-189
            check(vf_scene_set_height_f32(scene_, d, 2, 2));
        }
    }
    ~TerrainObject() { vf_scene_destroy(scene_); }
    TerrainObject(const TerrainObject&) = delete;
    TerrainObject& operator=(const TerrainObject&) = delete;

    void set_camera_look_at(std::tuple<float, float, float> eye, std::tuple<float, float, float> target,
                            std::tuple<float, float, float> up, float fovy_deg, float znear, float zfar) {
        const V3 e = v3(eye), t = v3(target), u = v3(up);
        vfcam::validate_camera_params(e, t, u, fovy_deg, znear, zfar);
        view_ = vfcam::look_at_rh(e, t, u);
        proj_ = vfcam::perspective_wgpu(vfcam::to_radians(fovy_deg), (float)width_ / (float)height_, znear, zfar);
        push_uniforms();
    }

    void set_height_from_r32f(py::object height_r32f) {
        // src/scene/mod.rs:229-231: PyReadonlyArray2<f32> extraction, then as_slice()
        if (!py::isinstance<py::array>(height_r32f)) throw py::type_error("argument 'height_r32f': expected a numpy float32 array");
        py::array arr = py::reinterpret_borrow<py::array>(height_r32f);
        if (arr.ndim() != 2 || !arr.dtype().is(py::dtype::of<float>()))
            throw py::type_error("argument 'height_r32f': expected a 2-D numpy array of dtype float32");
        if (!(arr.flags() & py::array::c_style)) throw std::runtime_error("height must be C-contiguous float32[H,W]");
        const uint32_t h = (uint32_t)arr.shape(0), w = (uint32_t)arr.shape(1);
        const float* p = static_cast<const float*>(arr.data());
        py::gil_scoped_release nogil;
        check(vf_scene_set_height_f32(scene_, p, w, h));
    }

    py::array_t<uint8_t> render_rgba() {
        auto out = pinned_array<uint8_t>({(py::ssize_t)height_, (py::ssize_t)width_, 4});
        uint8_t* p = out.mutable_data();
        encode_ms_ = 0.0f;
        {
            py::gil_scoped_release nogil;
            check(vf_scene_render_rgba(scene_, p));
        }
        return out;
    }

    // The frame is encoded on the device (vf_png_dev.cu) and only the file's bytes come back: src/scene/mod.rs:279-335
    // without the raw readback.  VF_PNG_HOST=1 keeps the round-1 path (raw readback + zlib on host threads).
    void render_png(const std::string& path) {
        if (std::getenv("VF_PNG_HOST")) {
            auto img = render_rgba();
            const uint8_t* p = img.data();
            py::gil_scoped_release nogil;
            const auto t0 = std::chrono::steady_clock::now();
            check(vf_png_write(path.c_str(), p, (int)width_, (int)height_));
            encode_ms_ = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
            return;
        }
        float u[44];
        check(vf_scene_get_uniforms(scene_, u));
        std::vector<std::string> paths{path};
        png_batch(u, 1, &paths, nullptr);
    }
    // n frames -> n .png files; cameras as for render_batch_rgba
    void render_batch_png(const std::vector<std::string>& paths, py::object cameras) {
        const std::vector<float> blocks = camera_blocks(cameras);
        const int n = (int)(blocks.size() / 44);
        if (n == 0) throw py::value_error("no cameras given");
        if ((size_t)n != paths.size()) throw py::value_error("need one path per camera");
        png_batch(blocks.data(), n, &paths, nullptr);
    }
    // n frames -> (pinned uint8 array (n, cap), uint32 sizes (n,)): file k is buf[k, :sizes[k]]; no copies on the host.
    // next_height: the float32[H,W] image (same size as the current one) the NEXT render call draws from; its upload overlaps
    // this batch (vf_scene_render_batch_png_next) -- the pipelined form of set_height_from_r32f + render_batch_png_raw.
    py::tuple render_batch_png_raw(py::object cameras, py::object next_height) {
        const std::vector<float> blocks = camera_blocks(cameras);
        const int n = (int)(blocks.size() / 44);
        if (n == 0) throw py::value_error("no cameras given");
        py::array next;
        if (!next_height.is_none()) {
            if (!py::isinstance<py::array>(next_height)) throw py::type_error("argument 'next_height': expected a numpy float32 array");
            next = py::reinterpret_borrow<py::array>(next_height);
            if (next.ndim() != 2 || !next.dtype().is(py::dtype::of<float>()))
                throw py::type_error("argument 'next_height': expected a 2-D numpy array of dtype float32");
            if (!(next.flags() & py::array::c_style)) throw std::runtime_error("height must be C-contiguous float32[H,W]");
        }
        py::array_t<uint32_t> sizes(n);
        uint32_t* sz = sizes.mutable_data();
        for (int attempt = 0; attempt < 2; ++attempt) {
            const size_t cap = png_slot_cap();
            auto buf = pinned_array<uint8_t>({(py::ssize_t)n, (py::ssize_t)cap});
            uint8_t* p = buf.mutable_data();
            bool ok;
            {
                py::gil_scoped_release nogil;
                if (!next_height.is_none())              // the repeat draws the same image again; the upload is already on its way
                    ok = png_render(blocks.data(), n, p, cap, sz, attempt == 0 ? static_cast<const float*>(next.data()) : nullptr,
                                    (uint32_t)next.shape(1), (uint32_t)next.shape(0), attempt == 0 ? 0 : VF_PNG_KEEP_IMAGE);
                else
                    ok = png_render(blocks.data(), n, p, cap, sz);
            }
            encode_ms_ = 0.0f;
            if (ok) {
                // every upload issued before this (synchronous) call has been consumed; the one staged by it is in flight
                pending_heights_.clear();
                if (!next_height.is_none()) pending_heights_.push_back(next);
                return py::make_tuple(buf, sizes);
            }
        }
        throw std::runtime_error("PNG batch did not fit its worst-case slots");
    }
    // n frames -> list of bytes objects (each a complete .png file)
    py::list render_batch_png_bytes(py::object cameras) {
        const std::vector<float> blocks = camera_blocks(cameras);
        const int n = (int)(blocks.size() / 44);
        if (n == 0) throw py::value_error("no cameras given");
        py::list out;
        png_batch(blocks.data(), n, nullptr, &out);
        return out;
    }

    py::array_t<float> debug_uniforms_f32() const {
        py::array_t<float> a(44);
        check(vf_scene_get_uniforms(scene_, a.mutable_data()));
        return a;
    }
    const char* debug_lut_format() const { return lut_unorm_ ? "Rgba8Unorm" : "Rgba8UnormSrgb"; }

    // ---- additive API (absent from the reference) ----
    void set_height_range(float mn, float mx) {
        if (!std::isfinite(mn) || !std::isfinite(mx)) throw py::value_error("min/max must be finite floats");
        if (mn >= mx) throw py::value_error("min must be < max");
        globals_.h_min = mn; globals_.h_max = mx;
        push_uniforms();
    }
    std::pair<float, float> set_auto_height_range() {
        float lo = 0, hi = 0;
        {
            py::gil_scoped_release nogil;
            check(vf_scene_auto_height_range(scene_, &lo, &hi));
        }
        globals_.h_min = lo; globals_.h_max = hi;
        push_uniforms();
        return {lo, hi};
    }
    void set_sun(float elevation_deg, float azimuth_deg) {
        if (!std::isfinite(elevation_deg) || !std::isfinite(azimuth_deg)) throw py::value_error("angles must be finite");
        globals_.sun_dir = sun_from_angles(elevation_deg, azimuth_deg);
        push_uniforms();
    }
    void set_exposure(float exposure) {
        if (!std::isfinite(exposure) || exposure <= 0.0f) throw py::value_error("exposure must be > 0");
        globals_.exposure = exposure;
        push_uniforms();
    }
    void set_sun_dir_exposure(V3 dir, float exposure) {
        globals_.sun_dir = dir; globals_.exposure = exposure;
        push_uniforms();
    }
    void set_spacing_exaggeration(float spacing, float exaggeration) {
        if (!(spacing > 0.0f) || !std::isfinite(spacing)) throw std::runtime_error("spacing components must be > 0");
        if (!(exaggeration > 0.0f) || !std::isfinite(exaggeration)) throw std::runtime_error("exaggeration must be > 0");
        globals_.spacing = spacing; globals_.exaggeration = exaggeration;
        push_uniforms();
    }
    void set_shading(const std::string& mode, bool depth) {
        if (mode == "shipped") mode_ = VF_MODE_SHIPPED;
        else if (mode == "northstar") mode_ = VF_MODE_NORTHSTAR;
        else throw py::value_error("mode must be 'shipped' or 'northstar'");
        depth_ = depth;
        push_uniforms();
    }
    void set_uniforms_raw(py::array_t<float, py::array::c_style | py::array::forcecast> u) {
        if (u.size() != 44) throw py::value_error("expected 44 float32 values");
        check(vf_scene_set_uniforms(scene_, u.data()));
    }

    // cameras: (N,44) float32 uniform blocks, or a sequence of (eye,target,up,fovy_deg,znear,zfar)
    std::vector<float> camera_blocks(py::object cameras) const {
        std::vector<float> blocks;
        if (py::isinstance<py::array>(cameras)) {
            auto a = py::array_t<float, py::array::c_style | py::array::forcecast>::ensure(cameras);
            if (!a || a.ndim() != 2 || a.shape(1) != 44) throw py::value_error("cameras array must have shape (N, 44)");
            blocks.assign(a.data(), a.data() + a.size());
            return blocks;
        }
        for (auto item : cameras) {
            auto t = item.cast<py::tuple>();
            if (t.size() != 6) throw py::value_error("camera tuples are (eye, target, up, fovy_deg, znear, zfar)");
            const V3 e = v3(t[0].cast<std::tuple<float, float, float>>()), tg = v3(t[1].cast<std::tuple<float, float, float>>()),
                     u = v3(t[2].cast<std::tuple<float, float, float>>());
            const float fovy = t[3].cast<float>(), zn = t[4].cast<float>(), zf = t[5].cast<float>();
            vfcam::validate_camera_params(e, tg, u, fovy, zn, zf);
            const M4 view = vfcam::look_at_rh(e, tg, u);
            const M4 proj = vfcam::perspective_wgpu(vfcam::to_radians(fovy), (float)width_ / (float)height_, zn, zf);
            float blk[44];
            globals_.to_uniforms(view, proj, blk);
            blocks.insert(blocks.end(), blk, blk + 44);
        }
        return blocks;
    }
    py::array_t<uint8_t> render_batch_rgba(py::object cameras) {
        const std::vector<float> blocks = camera_blocks(cameras);
        const int n = (int)(blocks.size() / 44);
        if (n == 0) throw py::value_error("no cameras given");
        auto out = pinned_array<uint8_t>({(py::ssize_t)n, (py::ssize_t)height_, (py::ssize_t)width_, 4});
        uint8_t* p = out.mutable_data();
        {
            py::gil_scoped_release nogil;
            check(vf_scene_render_batch(scene_, blocks.data(), n, p));
        }
        return out;
    }
    // ---- pipelined batches: upload k+1 and batch k overlap; see include/vf_cuda.h ----
    void set_height_from_r32f_async(py::object height_r32f) {
        if (!py::isinstance<py::array>(height_r32f)) throw py::type_error("argument 'height_r32f': expected a numpy float32 array");
        py::array arr = py::reinterpret_borrow<py::array>(height_r32f);
        if (arr.ndim() != 2 || !arr.dtype().is(py::dtype::of<float>()))
            throw py::type_error("argument 'height_r32f': expected a 2-D numpy array of dtype float32");
        if (!(arr.flags() & py::array::c_style)) throw std::runtime_error("height must be C-contiguous float32[H,W]");
        // The DMA is in flight after this call returns: the array stays referenced until the batch that
        // renders from it has been collected (or the scene is destroyed, which synchronises first).
        pending_heights_.push_back(arr);
        check(vf_scene_set_height_f32_async(scene_, static_cast<const float*>(arr.data()), (uint32_t)arr.shape(1), (uint32_t)arr.shape(0)));
    }
    int render_batch_rgba_async(py::object cameras) {
        const std::vector<float> blocks = camera_blocks(cameras);
        const int n = (int)(blocks.size() / 44);
        if (n == 0) throw py::value_error("no cameras given");
        auto out = pinned_array<uint8_t>({(py::ssize_t)n, (py::ssize_t)height_, (py::ssize_t)width_, 4});
        int ticket = -1;
        check(vf_scene_render_batch_async(scene_, blocks.data(), n, out.mutable_data(), &ticket));
        in_flight_[ticket] = out;
        ticket_heights_[ticket] = std::move(pending_heights_);   // uploads this batch (or an earlier one) consumes
        pending_heights_.clear();
        return ticket;
    }
    py::array_t<uint8_t> collect(int ticket) {
        auto it = in_flight_.find(ticket);
        if (it == in_flight_.end()) throw py::value_error("unknown ticket");
        {
            py::gil_scoped_release nogil;
            check(vf_scene_wait(scene_, ticket));
        }
        py::array_t<uint8_t> out = it->second;
        in_flight_.erase(it);
        ticket_heights_.erase(ticket);                        // the batch is done, so are the uploads queued before it
        return out;
    }
    // a frame left on the device by render_batch_device (frame k of a batch sits in slot k % 2)
    py::array_t<uint8_t> read_device_frame(int slot) {
        py::array_t<uint8_t> a({(py::ssize_t)height_, (py::ssize_t)width_, (py::ssize_t)4});
        check(vf_scene_read_frame(scene_, slot, a.mutable_data()));
        return a;
    }
    // frames stay on the device; returns elapsed milliseconds (host clock around a synchronous call)
    double render_batch_device(py::object cameras) {
        const std::vector<float> blocks = camera_blocks(cameras);
        const int n = (int)(blocks.size() / 44);
        if (n == 0) throw py::value_error("no cameras given");
        py::gil_scoped_release nogil;
        const auto t0 = std::chrono::steady_clock::now();
        check(vf_scene_render_device(scene_, blocks.data(), n));
        return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    }
    py::array_t<uint32_t> debug_visibility() {
        py::array_t<uint32_t> a({(py::ssize_t)height_, (py::ssize_t)width_});
        check(vf_scene_debug_visibility(scene_, a.mutable_data()));
        return a;
    }
    void enable_timing(bool on) { check(vf_scene_enable_timing(scene_, on ? 1 : 0)); }
    void set_stream(uintptr_t cuda_stream) { check(vf_scene_set_stream(scene_, reinterpret_cast<void*>(cuda_stream))); }
    void set_height_image(const vf_height* img) { check(vf_scene_set_height_image(scene_, img)); }
    // kernel statistics of the last frame plus the timing / traffic block the reference planned as render_metrics()
    // (ROADMAP.md:1959-1967: encode_ms, gpu_ms, readback_ms, total_ms, vram_bytes)
    py::dict render_metrics() const {
        vf_frame_stats_t st;
        check(vf_scene_last_stats(scene_, &st));
        vf_render_metrics_t rm;
        check(vf_scene_metrics(scene_, &rm));
        py::dict d;
        d["upload_ms"] = rm.upload_ms; d["gpu_ms"] = rm.gpu_ms; d["readback_ms"] = rm.readback_ms; d["encode_ms"] = encode_ms_;
        d["total_ms"] = rm.total_ms + encode_ms_; d["frames"] = rm.frames;
        d["h2d_bytes"] = rm.last_h2d_bytes; d["d2h_bytes"] = rm.last_d2h_bytes;
        d["h2d_bytes_total"] = rm.h2d_bytes; d["d2h_bytes_total"] = rm.d2h_bytes; d["vram_bytes"] = rm.vram_bytes;
        d["bin_pairs"] = st.bin_pairs; d["bin_capacity"] = st.bin_capacity;
        d["clipped_prims"] = st.clipped_prims; d["tiles"] = st.tiles; d["big_prims"] = st.big_prims; d["tma"] = st.tma;
        d["ms_geom"] = st.ms_geom; d["ms_bin"] = st.ms_bin; d["ms_raster"] = st.ms_raster; d["ms_total"] = st.ms_total;
        return d;
    }
    uint32_t width() const { return width_; }
    uint32_t height() const { return height_; }
    uint32_t grid() const { return grid_; }

private:
    // Host slot per file of a PNG batch.  The worst case (vf_png_file_bound, 1.13 x the raw frame) is rarely needed: the
    // slots start at a quarter of it and follow the largest file seen; a batch that does not fit is repeated with the
    // bound (png_render), so callers never see the difference.
    size_t png_slot_cap() const {
        const size_t bound = (vf_png_file_bound((int)width_, (int)height_) + 255) & ~(size_t)255;
        const size_t want = png_cap_hint_ ? png_cap_hint_ : std::max<size_t>(bound / 4, 1u << 16);
        return std::min(bound, (want + 65535) & ~(size_t)65535);
    }
    // renders n frames into `p` (n x cap bytes); returns false if a file did not fit (then nothing is valid)
    bool png_render(const float* blocks, int n, uint8_t* p, size_t cap, uint32_t* sizes, const float* next = nullptr, uint32_t next_w = 0,
                    uint32_t next_h = 0, int flags = 0) {
        const size_t bound = (vf_png_file_bound((int)width_, (int)height_) + 255) & ~(size_t)255;
        const int rc = (next || flags) ? vf_scene_render_batch_png_next(scene_, blocks, n, p, cap, sizes, next, next_w, next_h, flags)
                                       : vf_scene_render_batch_png(scene_, blocks, n, p, cap, sizes);
        if (rc == VF_E_INVALID && cap < bound && std::string(vf_last_error()).rfind("PNG file", 0) == 0) { png_cap_hint_ = bound; return false; }
        check(rc);
        uint32_t mx = 0;
        for (int k = 0; k < n; ++k) mx = std::max(mx, sizes[k]);
        png_cap_hint_ = std::max<size_t>(png_cap_hint_ == bound ? 0 : png_cap_hint_, (size_t)mx + mx / 4 + 4096);
        return true;
    }
    // render + encode on the device, files into pinned host memory, then to disk and / or bytes objects
    void png_batch(const float* blocks, int n, const std::vector<std::string>* paths, py::list* bytes_out) {
        size_t cap = png_slot_cap();
        auto buf = pinned_array<uint8_t>({(py::ssize_t)n, (py::ssize_t)cap});
        std::vector<uint32_t> sizes((size_t)n);
        uint8_t* p = buf.mutable_data();
        bool ok;
        { py::gil_scoped_release nogil; ok = png_render(blocks, n, p, cap, sizes.data()); }
        if (!ok) {                                               // (rare: the retry keeps the GIL)
            cap = png_slot_cap();
            buf = pinned_array<uint8_t>({(py::ssize_t)n, (py::ssize_t)cap});
            p = buf.mutable_data();
            if (!png_render(blocks, n, p, cap, sizes.data())) throw std::runtime_error("PNG batch did not fit its worst-case slots");
        }
        std::string io_error;
        {
            py::gil_scoped_release nogil;
            const auto t0 = std::chrono::steady_clock::now();
            if (paths) {
                for (int k = 0; k < n && io_error.empty(); ++k) {
                    FILE* f = std::fopen((*paths)[k].c_str(), "wb");
                    if (!f) { io_error = "cannot open '" + (*paths)[k] + "' for writing"; break; }
                    const bool ok = std::fwrite(p + (size_t)k * cap, 1, sizes[k], f) == sizes[k];
                    if (std::fclose(f) != 0 || !ok) io_error = "short write to '" + (*paths)[k] + "'";
                }
            }
            encode_ms_ = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();   // host share: file output
        }
        if (!io_error.empty()) throw std::runtime_error(io_error);
        if (bytes_out)
            for (int k = 0; k < n; ++k) bytes_out->append(py::bytes(reinterpret_cast<const char*>(p + (size_t)k * cap), sizes[k]));
    }
    void push_uniforms() {
        float u[44];
        globals_.to_uniforms(view_, proj_, u);
        check(vf_scene_set_uniforms(scene_, u));
        check(vf_scene_set_mode(scene_, mode_, depth_ ? 1 : 0, globals_.h_min, globals_.h_max));
    }
    uint32_t width_, height_, grid_ = 128;
    bool spike_, lut_unorm_ = false;
    vf_scene* scene_ = nullptr;
    std::vector<py::object> pending_heights_;
    std::map<int, std::vector<py::object>> ticket_heights_;
    std::map<int, py::array_t<uint8_t>> in_flight_;
    Globals globals_;
    M4 view_, proj_;
    int mode_ = VF_MODE_SHIPPED;
    bool depth_ = false;
    float encode_ms_ = 0.0f;
    size_t png_cap_hint_ = 0;
};

struct Scene : TerrainObject {
    Scene(uint32_t w, uint32_t h, py::object grid, py::object colormap) : TerrainObject(w, h, grid, colormap, false) {}
};
struct TerrainSpike : TerrainObject {
    TerrainSpike(uint32_t w, uint32_t h, py::object grid, py::object colormap) : TerrainObject(w, h, grid, colormap, true) {}
};

#include "module_renderer.inc"
