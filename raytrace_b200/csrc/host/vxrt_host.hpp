// vxrt_host.hpp — C++ host side above the C ABI (include/vxrt.h): the reference's `render`, `game` and `world`
// interfaces (src/render/mod.rs, src/game/mod.rs, src/world/chunk_storage.rs) with the same names and argument
// meaning, so that a caller shaped like src/bin/main.rs drives the CUDA path unchanged.  rustc is not in this
// image; INTEGRATION.md shows the Rust form of the same wiring.  No arithmetic of the hot path lives here.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <tuple>
#include <vector>

#include "vxrt.h"

namespace raytrace {

namespace constants {                         // src/render/constants.rs:9-33
constexpr int WINDOW_WIDTH = 1024, WINDOW_HEIGHT = 1024;
constexpr int BLUE_NOISE_WIDTH = 512, BLUE_NOISE_HEIGHT = 512, BLUE_NOISE_SIZE = 512 * 512 * 4;
constexpr int CHUNK_SIZE = 64, CHUNK_VOLUME = 64 * 64 * 64, ROOT_CHUNK_SIZE = 4, ROOT_BLOCK_SIZE = 256;
constexpr int SLICE_SIZE = 16;
}  // namespace constants

// The reference `.expect()`s every Vulkan call (panic); throwing keeps that caller-visible behaviour.
struct VxrtError : std::runtime_error {
    int code;
    VxrtError(int c, const std::string& what) : std::runtime_error("libvxrt error " + std::to_string(c) + ": " + what), code(c) {}
};
inline void check(const VxrtHandle* h, int rc) { if (rc != VXRT_OK) throw VxrtError(rc, vxrt_last_error(h)); }

// src/render/mod.rs:14-34
struct Camera {
    float origin[3] = {0.f, 0.f, 0.f};
    float heading = 3.14159265358979323846f * 0.5f;
    float pitch = 0.f;
};

// src/util.rs:9-22 (fp32; cgmath Vector3::cross)
struct TripleEulerVector { float forward[3], up[3], right[3]; };
inline TripleEulerVector compute_triple_euler_vector(float heading, float pitch) {
    const float half_pi = 3.14159265358979323846f / 2.f;
    const float ch = std::cos(heading), sh = std::sin(heading), cp = std::cos(pitch), sp = std::sin(pitch);
    const float cq = std::cos(pitch + half_pi), sq = std::sin(pitch + half_pi);
    TripleEulerVector t{{ch * cp, sh * cp, sp}, {ch * cq, sh * cq, sq}, {}};
    t.right[0] = t.forward[1] * t.up[2] - t.forward[2] * t.up[1];
    t.right[1] = t.forward[2] * t.up[0] - t.forward[0] * t.up[2];
    t.right[2] = t.forward[0] * t.up[1] - t.forward[1] * t.up[0];
    return t;
}

// src/world/chunk.rs:52-56
struct PackedChunkData {
    std::vector<uint32_t> materials;          // 64^3, x fastest (util.rs:94-106)
    std::vector<uint8_t> minefield;           // 64^3, or empty -> built on the GPU (k_build_minefield)
    PackedChunkData() : materials(constants::CHUNK_VOLUME, 0u) {}
};

// the part of src/world/chunk_storage.rs the renderer calls (:147-151); chunks never stored read as air
class ChunkStorage {
  public:
    using Coord = std::tuple<long, long, long>;
    PackedChunkData& borrow_packed_chunk_data(const Coord& c) { return chunks_[c]; }
    size_t len() const { return chunks_.size(); }
    // cut an [nz][ny][nx] window (texel = world + N/2) into chunks at coords -n/2 .. n/2-1
    void store_window(const uint32_t* words, int nx, int ny, int nz) {
        for (int cz = 0; cz < nz / 64; cz++) for (int cy = 0; cy < ny / 64; cy++) for (int cx = 0; cx < nx / 64; cx++) {
            PackedChunkData& d = chunks_[Coord(cx - nx / 128, cy - ny / 128, cz - nz / 128)];
            for (int z = 0; z < 64; z++) for (int y = 0; y < 64; y++)
                std::memcpy(&d.materials[(size_t)(z * 64 + y) * 64],
                            words + ((size_t)(cz * 64 + z) * ny + cy * 64 + y) * nx + cx * 64, 64 * sizeof(uint32_t));
        }
    }
  private:
    std::map<Coord, PackedChunkData> chunks_;
};

// src/game/mod.rs: the CALLER of the render boundary (headless: held controls are set by name)
class Game {
  public:
    Game() { camera_.origin[0] = -30.f; camera_.origin[1] = -128.f; camera_.origin[2] = 100.f; }   // mod.rs:52-56
    Game(float x, float y, float z, float heading, float pitch, float sun) : sun_angle_(sun) {      // mod.rs:44-51
        camera_.origin[0] = x; camera_.origin[1] = y; camera_.origin[2] = z; camera_.heading = heading; camera_.pitch = pitch;
    }
    void hold(const std::string& control, bool down) { held_[control] = down; }
    void tick(float dt) {                                                                            // mod.rs:61-96
        auto is_held = [&](const char* n) { auto it = held_.find(n); return it != held_.end() && it->second; };
        if (is_held("sunup")) sun_angle_ += dt; else if (is_held("sundown")) sun_angle_ -= dt;
        const float dx = is_held("left") ? -1.f : is_held("right") ? 1.f : 0.f;
        const float dy = is_held("backward") ? -1.f : is_held("forward") ? 1.f : 0.f;
        const float dz = is_held("down") ? -1.f : is_held("up") ? 1.f : 0.f;
        const float amount = dt * 50.f;
        TripleEulerVector t = compute_triple_euler_vector(camera_.heading, camera_.pitch);
        auto unit = [](float* v) { const float l = std::sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]); for (int i = 0; i < 3; i++) v[i] /= l; };
        unit(t.forward); unit(t.up); unit(t.right);
        for (int i = 0; i < 3; i++) camera_.origin[i] += amount * t.forward[i] * dy + amount * t.up[i] * dz + amount * t.right[i] * dx;
    }
    ChunkStorage& borrow_world_mut() { return world_; }
    const Camera& borrow_camera() const { return camera_; }
    Camera& borrow_camera_mut() { return camera_; }
    float get_sun_angle() const { return sun_angle_; }
  private:
    Camera camera_;
    ChunkStorage world_;
    std::map<std::string, bool> held_;
    float sun_angle_ = 0.f;
};


// src/render/pipeline/terrain_upload.rs: the sliding window.  Same positions, request queue and policy as the reference
// (Position :22-47, request_increase :292-322, request_decrease :324-349, request_move_towards :351-367, one 16-voxel slab per
// frame :277-286, render offset :29-36) with the window size taken from the handle.  A slab is gathered as ONE world box and
// rotated into texel order, texel_i = (world_i - origin_i * 64) mod N_i (the identity the reference's chunk-piece copies
// :84-211 implement); the toroidal scatter on the device is vxrt_upload_slice.
class TerrainUploadManager {
  public:
    struct Position {
        long origin[3]; int num_loaded_slices[3];
        void render_offset(const int root_chunks[3], int out[3]) const {                                  // :29-36
            for (int i = 0; i < 3; i++) out[i] = (int)((origin[i] + root_chunks[i] / 2) * constants::CHUNK_SIZE + num_loaded_slices[i] * constants::SLICE_SIZE);
        }
    };
    struct Request { long origin[3]; int num_slices[3]; int axis; Position new_position; };

    TerrainUploadManager(int nx, int ny, int nz) : window_{nx, ny, nz} {
        for (int i = 0; i < 3; i++) {
            root_chunks_[i] = window_[i] / constants::CHUNK_SIZE; slices_per_axis_[i] = window_[i] / constants::SLICE_SIZE;
            cpu_.origin[i] = gpu_.origin[i] = -(root_chunks_[i] / 2);                                       // Position::default, :39-46
            cpu_.num_loaded_slices[i] = gpu_.num_loaded_slices[i] = 0;
        }
    }
    void request_increase(int axis) {                                                                       // :292-322
        const Position old = cpu_;
        cpu_.num_loaded_slices[axis] += 1;
        if (cpu_.num_loaded_slices[axis] == slices_per_axis_[axis]) { cpu_.num_loaded_slices[axis] = 0; cpu_.origin[axis] += root_chunks_[axis]; }
        Request r;
        for (int i = 0; i < 3; i++) { r.origin[i] = old.origin[i]; r.num_slices[i] = old.num_loaded_slices[i]; }
        r.origin[axis] += root_chunks_[axis];
        r.axis = axis; r.new_position = cpu_;
        queue_.push_back(r);
    }
    void request_decrease(int axis) {                                                                       // :324-349
        if (cpu_.num_loaded_slices[axis] == 0) { cpu_.num_loaded_slices[axis] = slices_per_axis_[axis]; cpu_.origin[axis] -= root_chunks_[axis]; }
        cpu_.num_loaded_slices[axis] -= 1;
        Request r;
        for (int i = 0; i < 3; i++) { r.origin[i] = cpu_.origin[i]; r.num_slices[i] = cpu_.num_loaded_slices[i]; }
        r.axis = axis; r.new_position = cpu_;
        queue_.push_back(r);
    }
    void request_move_towards(const int desired_center[3]) {                                                // :351-367: one request, X before Y before Z
        int cur[3]; cpu_.render_offset(root_chunks_, cur);
        for (int axis = 0; axis < 3; axis++) {
            const int delta = desired_center[axis] - cur[axis];
            if (delta > constants::SLICE_SIZE) { request_increase(axis); return; }
            if (-delta > constants::SLICE_SIZE) { request_decrease(axis); return; }
        }
    }
    void get_render_offset(int out[3]) const { gpu_.render_offset(root_chunks_, out); }                    // :288-290
    size_t pending() const { return queue_.size(); }
    // :277-286 — consume at most one request: one slab per frame
    bool setup_next_request(ChunkStorage& world, VxrtHandle* h) {
        if (queue_.empty()) return false;
        const Request req = queue_.front();
        queue_.erase(queue_.begin());
        const int a = req.axis;
        long start[3]; int shape[3];
        for (int i = 0; i < 3; i++) { start[i] = req.origin[i] * constants::CHUNK_SIZE + (long)req.num_slices[i] * constants::SLICE_SIZE; shape[i] = window_[i]; }
        shape[a] = constants::SLICE_SIZE;
        const size_t n = (size_t)shape[0] * shape[1] * shape[2];
        slab_m_.assign(n, 0u); slab_f_.assign(n, 0);
        bool have_f = true;
        // slab texel (x, y, z) <- world voxel start + ((x, y, z) - 16 * num_slices) mod N on the off axes, start + (x, y, z) on `a`
        int shift[3];
        for (int i = 0; i < 3; i++) shift[i] = i == a ? 0 : (req.num_slices[i] * constants::SLICE_SIZE) % window_[i];
        auto fdiv64 = [](long v) { return v >= 0 ? v / 64 : -((-v + 63) / 64); };
        for (int z = 0; z < shape[2]; z++) for (int y = 0; y < shape[1]; y++) {
            const long wy = start[1] + ((y - shift[1]) % shape[1] + shape[1]) % shape[1], wz = start[2] + ((z - shift[2]) % shape[2] + shape[2]) % shape[2];
            const long cy = fdiv64(wy), cz = fdiv64(wz);
            for (int x = 0; x < shape[0];) {
                const long wx = start[0] + ((x - shift[0]) % shape[0] + shape[0]) % shape[0];
                const long cx = fdiv64(wx);
                // a run along x inside one chunk and before the rotation wraps
                int run = (int)((cx + 1) * 64 - wx);
                const int until_wrap = shift[0] > x ? shift[0] - x : shape[0] - x;
                if (run > until_wrap) run = until_wrap;
                if (run > shape[0] - x) run = shape[0] - x;
                PackedChunkData& c = world.borrow_packed_chunk_data(ChunkStorage::Coord(cx, cy, cz));
                const size_t src = (size_t)((wz - cz * 64) * 64 + (wy - cy * 64)) * 64 + (size_t)(wx - cx * 64);
                const size_t dst = ((size_t)z * shape[1] + y) * shape[0] + x;
                std::memcpy(&slab_m_[dst], &c.materials[src], (size_t)run * sizeof(uint32_t));
                if (c.minefield.empty()) have_f = false; else std::memcpy(&slab_f_[dst], &c.minefield[src], (size_t)run);
                x += run;
            }
        }
        const int offset = (req.num_slices[a] % slices_per_axis_[a]) * constants::SLICE_SIZE;             // :201-207
        check(h, vxrt_upload_slice(h, a, offset, slab_m_.data(), have_f ? slab_f_.data() : nullptr));
        gpu_ = req.new_position;
        return true;
    }
  private:
    int window_[3], root_chunks_[3], slices_per_axis_[3];
    Position cpu_, gpu_;
    std::vector<Request> queue_;
    std::vector<uint32_t> slab_m_;
    std::vector<uint8_t> slab_f_;
};

struct RenderOptions {                        // what constants.rs fixes at compile time
    int width = constants::WINDOW_WIDTH, height = constants::WINDOW_HEIGHT;
    int nx = constants::ROOT_BLOCK_SIZE, ny = constants::ROOT_BLOCK_SIZE, nz = constants::ROOT_BLOCK_SIZE;
    int mode = VXRT_MODE_FULL;
    uint32_t flags = VXRT_FLAG_DENOISE | VXRT_FLAG_FINALIZE;
    int device = 0;
    bool streaming = false;                   // pipeline.rs:174-189: the window follows the camera, one 16-voxel slab per frame
};

// src/render/pipeline/pipeline.rs: Pipeline::new (:36-76) + draw_frame (:134-255) over the ABI
class Pipeline {
  public:
    Pipeline(Game& game, const RenderOptions& o, const uint8_t* blue_noise /* 512*512*4 */) : opt_(o) {
        VxrtConfig cfg{};
        cfg.abi_version = VXRT_ABI_VERSION; cfg.width = o.width; cfg.height = o.height;
        cfg.nx = o.nx; cfg.ny = o.ny; cfg.nz = o.nz; cfg.mode = o.mode; cfg.flags = o.flags; cfg.device = o.device;
        check(nullptr, vxrt_create(&cfg, &h_));
        check(h_, vxrt_upload_blue_noise(h_, blue_noise, constants::BLUE_NOISE_SIZE));               // render_data.rs:110-133
        // render_data.rs:221-241: window chunk c holds world chunk c - n/2
        ChunkStorage& world = game.borrow_world_mut();
        for (int cz = 0; cz < o.nz / 64; cz++) for (int cy = 0; cy < o.ny / 64; cy++) for (int cx = 0; cx < o.nx / 64; cx++) {
            PackedChunkData& c = world.borrow_packed_chunk_data(ChunkStorage::Coord(cx - o.nx / 128, cy - o.ny / 128, cz - o.nz / 128));
            check(h_, vxrt_upload_chunk(h_, cx, cy, cz, c.materials.data(), c.minefield.empty() ? nullptr : c.minefield.data()));
        }
        if (o.streaming) tum_.reset(new TerrainUploadManager(o.nx, o.ny, o.nz));
        frame_.resize((size_t)o.width * o.height * 4);
        check(h_, vxrt_host_register(h_, frame_.data(), frame_.size()));
        std::memset(&uniform_data_, 0, sizeof uniform_data_);
    }
    ~Pipeline() { if (h_) { vxrt_host_unregister(h_, frame_.data()); vxrt_destroy(h_); } }
    Pipeline(const Pipeline&) = delete;
    Pipeline& operator=(const Pipeline&) = delete;

    // pipeline.rs:191-235: fill the uniform block, submit, "present" (= read back into frame())
    void draw_frame(Game& game) {
        const Camera& cam = game.borrow_camera();
        if (tum_) {                                                                                  // pipeline.rs:174-189 (y pinned to 0 there)
            const int want[3] = {(int)cam.origin[0], 0, (int)cam.origin[2]};
            tum_->request_move_towards(want);
            tum_->setup_next_request(game.borrow_world_mut(), h_);
            tum_->get_render_offset(render_offset_);
        }
        const TripleEulerVector t = compute_triple_euler_vector(cam.heading, cam.pitch);
        VxrtFrameParams& u = uniform_data_;
        for (int i = 0; i < 3; i++) {
            u.origin[i] = cam.origin[i]; u.forward[i] = t.forward[i]; u.up[i] = t.up[i] * 0.4f; u.right[i] = t.right[i] * 0.4f;
            u.rotation[i] = render_offset_[i]; u.space_offset[i] = render_offset_[i];
        }
        u.seed = (u.seed + 1) % (uint32_t)constants::BLUE_NOISE_SIZE;                                // pipeline.rs:201
        u.sun_angle = game.get_sun_angle();
        check(h_, vxrt_set_frame(h_, &u));
        check(h_, vxrt_render_frame(h_));
        check(h_, vxrt_read_output(h_, frame_.data(), frame_.size()));
        // pipeline.rs:213-227: set after the submit, so it describes this frame to the next one (read by VXRT_FLAG_REPROJECT)
        for (int i = 0; i < 3; i++) u.old_origin[i] = u.origin[i];
        {   // cgmath 0.17.0 Matrix3::invert of from_cols(right*0.4, up*0.4, forward): cofactor rows / determinant, transposed
            const float* a = u.right; const float* b = u.up; const float* c = u.forward;
            auto cross = [](const float* p, const float* q, float* o) { o[0] = p[1] * q[2] - p[2] * q[1]; o[1] = p[2] * q[0] - p[0] * q[2]; o[2] = p[0] * q[1] - p[1] * q[0]; };
            const float det = a[0] * (b[1] * c[2] - c[1] * b[2]) - b[0] * (a[1] * c[2] - c[1] * a[2]) + c[0] * (a[1] * b[2] - b[1] * a[2]);
            float r0[3] = {0, 0, 0}, r1[3] = {0, 0, 0}, r2[3] = {0, 0, 0};
            if (det != 0.0f) { cross(b, c, r0); cross(c, a, r1); cross(a, b, r2); for (int i = 0; i < 3; i++) { r0[i] /= det; r1[i] /= det; r2[i] /= det; } }
            for (int k = 0; k < 3; k++) {
                float* col = k == 0 ? u.old_transform_c0 : k == 1 ? u.old_transform_c1 : u.old_transform_c2;
                col[0] = r0[k]; col[1] = r1[k]; col[2] = r2[k];
            }
        }
    }
    void set_render_offset(int x, int y, int z) { render_offset_[0] = x; render_offset_[1] = y; render_offset_[2] = z; }
    const std::vector<uint8_t>& frame() const { return frame_; }
    const VxrtFrameParams& uniform_data() const { return uniform_data_; }
    VxrtFrameStats stats() { VxrtFrameStats s{}; check(h_, vxrt_frame_stats(h_, &s)); return s; }
    VxrtHandle* handle() { return h_; }
    const RenderOptions& options() const { return opt_; }
    const int* render_offset() const { return render_offset_; }
  private:
    RenderOptions opt_;
    VxrtHandle* h_ = nullptr;
    VxrtFrameParams uniform_data_;
    int render_offset_[3] = {0, 0, 0};
    std::unique_ptr<TerrainUploadManager> tum_;
    std::vector<uint8_t> frame_;
};

// src/render/mod.rs:36-43 (no window / event loop: headless)
inline std::unique_ptr<Pipeline> create_instance(Game& game, const RenderOptions& o, const uint8_t* blue_noise) {
    return std::unique_ptr<Pipeline>(new Pipeline(game, o, blue_noise));
}

}  // namespace raytrace
