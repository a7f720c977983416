// ORACLE — TEST INFRASTRUCTURE ONLY (see ref_math.hpp header).  PARITY UNPINNED.
//
// Follows src/main.rs:219-343 (apply_box_filter*: three 3x3 cross-weighted filters) and
// src/main.rs:374-437 (the seven display modes of draw_to_sdl), operating on the Vec<Pixel> the
// render threads fill.  The window clipping at the image border is the reference's
// (main.rs:324-343); like the reference's loops it assumes width, height >= 3.
#pragma once
#include <cmath>
#include <vector>

#include "../include/ottomarcher.h"
#include "ref_render.hpp"

namespace ref {

constexpr float SQRT2_INV = 0.7071067811865475244f;  // utils.rs:31

inline Vec3 px_sum(const om_pixel& p) { return Vec3(p.sum[0], p.sum[1], p.sum[2]); }
inline void put_rgb(uint8_t* rgb, size_t idx, const uint8_t c[3]) { rgb[3 * idx] = c[0]; rgb[3 * idx + 1] = c[1]; rgb[3 * idx + 2] = c[2]; }

// apply_box_filter_ij_{samples,depth,id}  main.rs:222-318
inline void box_filter_ij(int mode, const om_pixel* px, int W, uint8_t* rgb, int i, int j, int min_x, int max_x, int min_y, int max_y) {
    float total_weight = 0.f;
    Vec3 color(0.f, 0.f, 0.f);
    const size_t aux = size_t(i) + size_t(j) * size_t(W);
    const float di = px[aux].avg_depth;
    const uint64_t state = px[aux].bloom;
    if (mode == 1 && std::isinf(di)) { put_rgb(rgb, aux, px[aux].color); return; }  // main.rs:252-259
    for (int y = min_y; y <= max_y; y++) {
        for (int x = min_x; x <= max_x; x++) {
            const om_pixel& q = px[size_t(i + x) + size_t(j + y) * size_t(W)];
            const float is_diagonal = (x != 0 && y != 0) ? 1.f : 0.f;
            if (mode == 0) {  // main.rs:231-236
                const float n = (float)q.n;
                const float diag_w = 1.f - (1.f - SQRT2_INV) * is_diagonal;
                total_weight += n * diag_w;
                color = color + px_sum(q) * diag_w;  // AddAssign vec3.rs = componentwise +
            } else if (mode == 1) {  // main.rs:264-274
                const float d = q.avg_depth;
                const float w = 1.f / (1.f + std::fabs(d - di));
                const float diag_w = w * (1.f - (1.f - SQRT2_INV) * is_diagonal);
                total_weight += diag_w;
                const Vec3 c = px_sum(q) / (float)q.n;
                color = color + diag_w * c;
            } else {  // main.rs:297-306
                const float same_value = (q.bloom == state) ? 1.f : 0.f;
                const float partial_value = ((q.bloom & state) == state) ? 1.f : 0.f;
                const float w = same_value + partial_value;
                const float diag_w = w * (1.f - (1.f - SQRT2_INV) * is_diagonal);
                total_weight += diag_w;
                const Vec3 c = px_sum(q) / (float)q.n;
                color = color + diag_w * c;
            }
        }
    }
    uint8_t c[3];
    to_u8x3(normalize_color(color / total_weight), c);
    put_rgb(rgb, aux, c);
}

// apply_box_filter main.rs:321-343
inline void box_filter(int mode, const om_pixel* px, int H, int W, uint8_t* rgb) {
    for (int j = 1; j < H - 1; j++) {
        for (int i = 1; i < W - 1; i++) {
            if (j == 1) {
                box_filter_ij(mode, px, W, rgb, i, 0, -1, 1, 0, 1);
                box_filter_ij(mode, px, W, rgb, i, H - 1, -1, 1, -1, 0);
            }
            box_filter_ij(mode, px, W, rgb, i, j, -1, 1, -1, 1);
        }
        box_filter_ij(mode, px, W, rgb, 0, j, 0, 1, -1, 1);
        box_filter_ij(mode, px, W, rgb, W - 1, j, -1, 0, -1, 1);
    }
    box_filter_ij(mode, px, W, rgb, 0, 0, 0, 1, 0, 1);
    box_filter_ij(mode, px, W, rgb, W - 1, 0, -1, 0, 0, 1);
    box_filter_ij(mode, px, W, rgb, 0, H - 1, 0, 1, -1, 0);
    box_filter_ij(mode, px, W, rgb, W - 1, H - 1, -1, 0, -1, 0);
}

// one refresh of draw_to_sdl's pixel buffer, main.rs:374-437.  mode: 0 normal, 1 samples, 2 sample-weighted blur,
// 3 depth, 4 depth-weighted blur, 5 ids, 6 id-weighted blur (main.rs:361-368)
inline void view(int mode, const om_pixel* px, int W, int H, uint8_t* rgb) {
    const size_t n = size_t(W) * size_t(H);
    if (mode == 0) {
        for (size_t pos = 0; pos < n; pos++) put_rgb(rgb, pos, px[pos].color);
    } else if (mode == 1) {
        uint32_t max_samples = 1;
        for (size_t pos = 0; pos < n; pos++) if (px[pos].n > max_samples) max_samples = px[pos].n;
        for (size_t pos = 0; pos < n; pos++) {
            const float smpls = (float)px[pos].n / (float)max_samples;
            uint8_t c[3];
            to_u8x3(normalize_color(Vec3(smpls, smpls, smpls)), c);
            put_rgb(rgb, pos, c);
        }
    } else if (mode == 3) {
        float max_depth = -1.f;
        for (size_t pos = 0; pos < n; pos++) { const float d = px[pos].avg_depth; if (d > max_depth && !std::isinf(d)) max_depth = d; }
        for (size_t pos = 0; pos < n; pos++) {
            const float depth_0to1 = px[pos].avg_depth / max_depth;
            const bool is_inf = std::isinf(depth_0to1);
            const float depth_rb = is_inf ? 0.f : depth_0to1, depth_g = is_inf ? 1.f : depth_0to1;
            uint8_t c[3];
            to_u8x3(normalize_color(Vec3(depth_rb, depth_g, depth_rb)), c);
            put_rgb(rgb, pos, c);
        }
    } else if (mode == 2) box_filter(0, px, H, W, rgb);
    else if (mode == 4) box_filter(1, px, H, W, rgb);
    else if (mode == 6) box_filter(2, px, H, W, rgb);
    else if (mode == 5) {
        for (size_t pos = 0; pos < n; pos++) { uint8_t c[3]; u64_to_color(scramble(px[pos].bloom), c); put_rgb(rgb, pos, c); }
    }
}

}  // namespace ref
