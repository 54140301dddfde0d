// Headless rasteriser: what Interface::render draws (SimulatorCPP/Interface.cpp:55-135 -- clear, render_grid :175-220,
// render_boxels :411-517, render_edges :137-173) into an RGB8 buffer, as a CUDA kernel instead of fixed-function OpenGL.
// The reference's GL path itself is unavailable on this box (no graphics stack); there are no shaders in it, only
// glColor / GL_QUADS / GL_LINES calls, whose coverage and colour rules are restated here:
//
//   camera      Camera::world_to_camera (Camera.cpp:66-70): ndc = (world - pos) * (2 / size), y negated for image cameras, so
//               that row 0 of the glReadPixels buffer is the top of the view.  Pixel (c, r) has its centre at window
//               coordinates (c + 0.5, r + 0.5).
//   background  (244, 245, 247) / 255, white with hide_background.
//   grid        lines at every multiple of 0.1 (one voxel), colour 0.8 grey.
//   voxels      one quad per boxel (TL, TR, BR, BL), objects and boxels in load order (later ones paint over earlier ones):
//               FIXED / RIGID 0.15 grey, SOFT 0.75 grey, actuators by their current length ratio through the two
//               colour ramps of Colors.h (sampled ColorBrewer "Oranges" for horizontal, "Blues" for vertical actuators:
//               regenerated here from the nine published anchors of each ramp).
//   edges       surface edges as one-pixel lines, colour (0, 63, 92) / 255.
//
// One thread per pixel, one CTA per 16 x 16 tile; the primitives of the image are staged through shared memory in chunks and
// culled against the tile first.  Coverage: a pixel belongs to a quad when its centre lies in one of the two triangles
// (TL, TR, BR) / (TL, BR, BL); to a line when its centre is within half a pixel of the segment.  Pixel-exact agreement with
// a particular GL implementation is not claimed (GL leaves line rasterisation and shared-edge ties to the implementation).
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <stdexcept>
#include <string>
#include <vector>

#include "engine.h"

namespace mevo {

#define CUDA_CHECK(expr)                                                                          \
    do {                                                                                          \
        cudaError_t _e = (expr);                                                                  \
        if (_e != cudaSuccess)                                                                    \
            throw std::runtime_error(std::string("CUDA error: ") + cudaGetErrorString(_e) + " at " + \
                                     __FILE__ + ":" + std::to_string(__LINE__) + " (" #expr ")");   \
    } while (0)

namespace {

struct Prim {                 // 48 bytes: a quad (4 corners) or a line (2 end points), pixel coordinates
    float x[4], y[4];
    float bx0, by0, bx1, by1; // bounding box (grown by half a pixel for lines)
    unsigned rgb_kind;        // bits 0-23 colour (r | g << 8 | b << 16), bit 31: line
    unsigned pad[3];
};

struct RenderParams {
    const Prim* prims;
    int n_prims;
    int w, h;
    // grid lines: pixel x of world x = 0 and pixel spacing of one voxel (0.1), same for y; valid index ranges
    float gx0, gdx, gy0, gdy;
    int gi_min, gi_max, gj_min, gj_max;
    unsigned background, grid_rgb;
    int draw_grid, clear;
    unsigned char* out;       // h x w x 3
};

__device__ __forceinline__ float edge_fn(float ax, float ay, float bx, float by, float px, float py) {
    return (bx - ax) * (py - ay) - (by - ay) * (px - ax);
}

__device__ __forceinline__ bool in_triangle(float ax, float ay, float bx, float by, float cx, float cy, float px, float py) {
    const float e0 = edge_fn(ax, ay, bx, by, px, py), e1 = edge_fn(bx, by, cx, cy, px, py), e2 = edge_fn(cx, cy, ax, ay, px, py);
    return (e0 >= 0.f && e1 >= 0.f && e2 >= 0.f) || (e0 <= 0.f && e1 <= 0.f && e2 <= 0.f);
}

constexpr int kTile = 16;
constexpr int kChunk = 128;

__global__ void __launch_bounds__(kTile* kTile) render_kernel(RenderParams p) {
    __shared__ Prim sp[kChunk];
    __shared__ int keep[kChunk];
    const int c = blockIdx.x * kTile + threadIdx.x, r = blockIdx.y * kTile + threadIdx.y;
    const int tid = threadIdx.y * kTile + threadIdx.x;
    const float px = c + 0.5f, py = r + 0.5f;
    const float tx0 = blockIdx.x * kTile, ty0 = blockIdx.y * kTile, tx1 = tx0 + kTile, ty1 = ty0 + kTile;
    const bool inside = c < p.w && r < p.h;
    unsigned colour = p.background;
    bool written = p.clear != 0;
    if (p.draw_grid) {
        // nearest grid line in x and in y; a line covers the pixel whose centre is within half a pixel of it
        const float fi = (px - p.gx0) / p.gdx, fj = (py - p.gy0) / p.gdy;
        const int i = __float2int_rn(fi), j = __float2int_rn(fj);
        const bool on_x = i >= p.gi_min && i <= p.gi_max && fabsf(px - (p.gx0 + i * p.gdx)) <= 0.5f;
        const bool on_y = j >= p.gj_min && j <= p.gj_max && fabsf(py - (p.gy0 + j * p.gdy)) <= 0.5f;
        if (on_x || on_y) { colour = p.grid_rgb; written = true; }
    }
    for (int base = 0; base < p.n_prims; base += kChunk) {
        const int m = min(kChunk, p.n_prims - base);
        __syncthreads();
        if (tid < m) {
            const Prim q = p.prims[base + tid];
            sp[tid] = q;
            keep[tid] = !(q.bx1 < tx0 || q.bx0 > tx1 || q.by1 < ty0 || q.by0 > ty1);
        }
        __syncthreads();
        for (int k = 0; k < m; k++) {                       // painter's order
            if (!keep[k]) continue;
            const Prim& q = sp[k];
            bool hit;
            if (q.rgb_kind >> 31) {
                const float dx = q.x[1] - q.x[0], dy = q.y[1] - q.y[0];
                const float len2 = dx * dx + dy * dy;
                float t = len2 > 0.f ? ((px - q.x[0]) * dx + (py - q.y[0]) * dy) / len2 : 0.f;
                t = fminf(fmaxf(t, 0.f), 1.f);
                const float ex = px - (q.x[0] + t * dx), ey = py - (q.y[0] + t * dy);
                hit = ex * ex + ey * ey <= 0.25f;
            } else {
                hit = px >= q.bx0 && px <= q.bx1 && py >= q.by0 && py <= q.by1 &&
                      (in_triangle(q.x[0], q.y[0], q.x[1], q.y[1], q.x[2], q.y[2], px, py) ||
                       in_triangle(q.x[0], q.y[0], q.x[2], q.y[2], q.x[3], q.y[3], px, py));
            }
            if (hit) { colour = q.rgb_kind & 0xffffffu; written = true; }
        }
        __syncthreads();
    }
    if (inside && written) {
        unsigned char* o = p.out + ((size_t)r * p.w + c) * 3;
        o[0] = colour & 0xff; o[1] = (colour >> 8) & 0xff; o[2] = (colour >> 16) & 0xff;
    }
}

unsigned pack_rgb(double r, double g, double b) {
    auto q = [](double v) { return (unsigned)std::lround(std::min(1.0, std::max(0.0, v)) * 255.0); };   // GL float -> ubyte
    return q(r) | (q(g) << 8) | (q(b) << 16);
}

// ColorBrewer 9-class sequential ramps (the anchors matplotlib's "Oranges" / "Blues" interpolate); Colors.h holds them
// sampled at 100 points: row i = ramp(i / 99).
const double kOranges[9][3] = {{255, 245, 235}, {254, 230, 206}, {253, 208, 162}, {253, 174, 107}, {253, 141, 60},
                               {241, 105, 19},  {217, 72, 1},    {166, 54, 3},    {127, 39, 4}};
const double kBlues[9][3] = {{247, 251, 255}, {222, 235, 247}, {198, 219, 239}, {158, 202, 225}, {107, 174, 214},
                             {66, 146, 198},  {33, 113, 181},  {8, 81, 156},    {8, 48, 107}};

unsigned ramp_row(const double (*anchors)[3], int row) {
    const double t = std::min(99, std::max(0, row)) / 99.0 * 8.0;
    const int k = std::min((int)t, 7);
    const double f = t - k;
    double c[3];
    for (int i = 0; i < 3; i++) c[i] = (anchors[k][i] * (1 - f) + anchors[k + 1][i] * f) / 255.0;
    return pack_rgb(c[0], c[1], c[2]);
}

// setColorInterval(actuation, is_act_vert) (Interface.cpp:378-409)
unsigned actuator_colour(double actuation, bool vertical) {
    const double (*ramp)[3] = vertical ? kBlues : kOranges;
    int index;
    if (actuation > 1) {
        index = (int)((2.0 - actuation) * 50.0);
        index = std::min(50, std::max(0, index));
        return ramp_row(ramp, index);
    }
    index = (int)((1.0 - actuation) * 50.0);              // actuation == 1 keeps GL's previous colour in the reference; row 50 here
    index = std::min(49, std::max(0, index));             // (the reference's clamp to 50 would read past the table's last row)
    return ramp_row(ramp, index + 50);
}

}  // namespace

void Batch::render_device(int env, double cam_x, double cam_y, double cam_w, double cam_h, int res_w, int res_h, int flags,
                          unsigned char* out_device) {
    CUDA_CHECK(cudaSetDevice(device_));
    if (!finalized_) finalize();
    if (res_w <= 0 || res_h <= 0 || res_w > 16384 || res_h > 16384) throw std::runtime_error("render: bad resolution");
    if (!(cam_w > 0) || !(cam_h > 0)) throw std::runtime_error("render: camera size must be positive");
    const bool hide_background = flags & 1, hide_grid = flags & 2, hide_edges = flags & 4, hide_boxels = flags & 8, dont_clear = flags & 16;
    // current positions in the reference's numbering (downloads the environment if the device copy is newer)
    const int P = world(env).num_points();
    std::vector<double> pos(2 * (size_t)P);
    get_state(env, pos.data(), nullptr, nullptr, nullptr);
    const World& W = world(env);
    const double sx = 2.0 / cam_w, sy = 2.0 / cam_h;
    auto to_px = [&](int pt, float& X, float& Y) {
        X = (float)(((pos[2 * pt] - cam_x) * sx + 1.0) * res_w * 0.5);
        Y = (float)((-(pos[2 * pt + 1] - cam_y) * sy + 1.0) * res_h * 0.5);
    };
    auto dist = [&](int a, int b) { return std::hypot(pos[2 * a] - pos[2 * b], pos[2 * a + 1] - pos[2 * b + 1]); };
    std::vector<Prim> prims;
    auto finish = [&](Prim& q, int n, float grow) {
        q.bx0 = q.bx1 = q.x[0]; q.by0 = q.by1 = q.y[0];
        for (int k = 1; k < n; k++) {
            q.bx0 = std::min(q.bx0, q.x[k]); q.bx1 = std::max(q.bx1, q.x[k]);
            q.by0 = std::min(q.by0, q.y[k]); q.by1 = std::max(q.by1, q.y[k]);
        }
        q.bx0 -= grow; q.by0 -= grow; q.bx1 += grow; q.by1 += grow;
        prims.push_back(q);
    };
    if (!hide_boxels) {
        for (const VoxelObject& ob : W.objects)
            for (const Boxel& b : ob.boxels) {
                Prim q = {};
                unsigned rgb;
                if (b.cell_type == CELL_SOFT) rgb = pack_rgb(0.75, 0.75, 0.75);
                else if (b.cell_type == CELL_ACT_H || b.cell_type == CELL_ACT_V) {
                    const bool vert = b.cell_type == CELL_ACT_V;
                    const int e1 = vert ? b.ed[3] : b.ed[0], e2 = vert ? b.ed[1] : b.ed[2];     // left,right | top,bot
                    const double avg = (dist(W.ea[e1], W.eb[e1]) + dist(W.ea[e2], W.eb[e2])) / 2.0;
                    rgb = actuator_colour(avg / W.el0[e1], vert);
                } else rgb = pack_rgb(0.15, 0.15, 0.15);
                for (int k = 0; k < 4; k++) to_px(b.pt[k], q.x[k], q.y[k]);
                q.rgb_kind = rgb;
                finish(q, 4, 0.f);
            }
    }
    if (!hide_edges) {
        const unsigned rgb = pack_rgb(0.0, 63.0 / 255.0, 92.0 / 255.0);
        for (int e = 0; e < W.num_edges(); e++) {
            if (!W.on_surface[e]) continue;
            Prim q = {};
            to_px(W.ea[e], q.x[0], q.y[0]);
            to_px(W.eb[e], q.x[1], q.y[1]);
            q.rgb_kind = rgb | (1u << 31);
            finish(q, 2, 0.5f);
        }
    }
    RenderParams p = {};
    Prim* d_prims = nullptr;
    CUDA_CHECK(cudaMalloc(&d_prims, sizeof(Prim) * std::max<size_t>(1, prims.size())));
    try {
        if (!prims.empty())
            CUDA_CHECK(cudaMemcpyAsync(d_prims, prims.data(), sizeof(Prim) * prims.size(), cudaMemcpyHostToDevice, (cudaStream_t)stream_));
        p.prims = d_prims; p.n_prims = (int)prims.size();
        p.w = res_w; p.h = res_h;
        p.gdx = (float)(0.1 * sx * res_w * 0.5); p.gx0 = (float)(((0.0 - cam_x) * sx + 1.0) * res_w * 0.5);
        p.gdy = (float)(-0.1 * sy * res_h * 0.5); p.gy0 = (float)((-(0.0 - cam_y) * sy + 1.0) * res_h * 0.5);
        // render_grid's index ranges (Interface.cpp:177-180; note it uses the full size, not half, on each side)
        p.gi_min = (int)((cam_x - cam_w) * 10.0) - 1; p.gi_max = (int)((cam_x + cam_w) * 10.0) + 1;
        p.gj_min = (int)((cam_y - cam_h) * 10.0) - 1; p.gj_max = (int)((cam_y + cam_h) * 10.0) + 1;
        p.background = hide_background ? pack_rgb(1, 1, 1) : pack_rgb(244.0 / 255.0, 245.0 / 255.0, 247.0 / 255.0);
        p.grid_rgb = pack_rgb(0.8, 0.8, 0.8);
        p.draw_grid = hide_grid ? 0 : 1;
        p.clear = dont_clear ? 0 : 1;
        p.out = out_device;
        const dim3 grid((res_w + kTile - 1) / kTile, (res_h + kTile - 1) / kTile), block(kTile, kTile);
        render_kernel<<<grid, block, 0, (cudaStream_t)stream_>>>(p);
        CUDA_CHECK(cudaGetLastError());
        CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream_));      // d_prims is freed below
    } catch (...) { cudaFree(d_prims); throw; }
    cudaFree(d_prims);
}

void Batch::render(int env, double cam_x, double cam_y, double cam_w, double cam_h, int res_w, int res_h, int flags,
                   unsigned char* out_host) {
    CUDA_CHECK(cudaSetDevice(device_));
    const size_t bytes = (size_t)res_w * res_h * 3;
    unsigned char* d = nullptr;
    CUDA_CHECK(cudaMalloc(&d, bytes));
    try {
        if (flags & 16) CUDA_CHECK(cudaMemcpy(d, out_host, bytes, cudaMemcpyHostToDevice));   // dont_clear: paint over the caller's image
        render_device(env, cam_x, cam_y, cam_w, cam_h, res_w, res_h, flags, d);
        CUDA_CHECK(cudaMemcpy(out_host, d, bytes, cudaMemcpyDeviceToHost));
    } catch (...) { cudaFree(d); throw; }
    cudaFree(d);
}

}  // namespace mevo
