// capi_client.cpp -- a plain C++ host of include/sr_capi.h (no OpenCV, no Python): the same calls the
// shim headers make, checked against the CPU oracle linked into this test binary.
// usage: capi_client            -> runs the checks, exit code 0 = parity
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include "../../include/sr_capi.h"
#include "../../oracle/sr_oracle.h"

#define CHECK(c) do { int rc_ = (c); if (rc_ != 0) { fprintf(stderr, "%s failed: %d (%s)\n", #c, rc_, sr_last_error(ctx)); return 1; } } while (0)

int main() {
    sr_ctx* ctx = nullptr;
    if (sr_ctx_create(0, &ctx) != SR_OK) { fprintf(stderr, "no context: %s\n", sr_last_error(nullptr)); return 2; }
    const int w = 420, h = 96, D = 56;
    std::vector<uint8_t> L((size_t)w * h * 3), R((size_t)w * h * 3);
    orc_synth_pair(0xB200, 1, w, h, 0, D, 0, L.data(), R.data());
    sr_image li{L.data(), w, h, (size_t)w * 3}, ri{R.data(), w, h, (size_t)w * 3};
    int bad = 0;
    // BlockSearch(left,right,17,0,D).computeDisparityMapRight(0.9) and (9, smooth 1.0)
    for (int cfg = 0; cfg < 2; cfg++) {
        int block = cfg ? 9 : 17;
        double smooth = cfg ? 1.0 : 0.9;
        sr_search_params p{SR_VARIANT_RIGHT, block, 0, D, smooth, 0, 19.0, SR_COST_SSD_L2, 0, 0};
        std::vector<double> got((size_t)w * h), want((size_t)w * h);
        int mb = 0;
        CHECK(sr_block_search(ctx, &li, &ri, &p, got.data(), (size_t)w * 8, &mb));
        if (orc_block_search_right_fast(L.data(), w, h, (size_t)w * 3, R.data(), w, h, (size_t)w * 3, block, 0, D, smooth, ORC_COST_SSD_L2, 0, h, want.data(), 0) != ORC_OK) return 3;
        size_t diff = 0;
        for (size_t i = 0; i < got.size(); i++) diff += got[i] != want[i];
        printf("right block %d smooth %.1f: %zu mismatching pixels, max block %d, kernel %s\n", block, smooth, diff, mb, sr_last_kernel(ctx));
        bad += diff != 0 || mb != block;
        // main.cpp:41-65 on that map: u8 hand-off, outliers, depth, vertices, faces
        if (cfg == 0) {
            std::vector<float> d32((size_t)w * h), depth_g((size_t)w * h), depth_c((size_t)w * h);
            orc_quantize_u8(want.data(), w * h, d32.data());
            std::vector<float> dc = d32;
            CHECK(sr_remove_outliers(ctx, d32.data(), w, h, (size_t)w * 4, 21, 1.5f, 0.8f));
            orc_remove_outliers(dc.data(), w, h, 21, 1.5f, 0.8f);
            bad += memcmp(d32.data(), dc.data(), d32.size() * 4) != 0;
            CHECK(sr_disparity_to_depth(ctx, d32.data(), depth_g.data(), w, h, 700.f, 1.f));
            orc_disparity_to_depth(dc.data(), depth_c.data(), w, h, 700.f, 1.f);
            bad += memcmp(depth_g.data(), depth_c.data(), depth_g.size() * 4) != 0;
            std::vector<sr_vertex> vg((size_t)w * h);
            std::vector<orc_vertex> vc((size_t)w * h);
            CHECK(sr_backproject(ctx, R.data(), (size_t)w * 3, depth_g.data(), w, h, 700.f, 700.f, w / 2.f, h / 2.f, vg.data()));
            orc_backproject(R.data(), (size_t)w * 3, depth_c.data(), w, h, 700.f, 700.f, w / 2.f, h / 2.f, vc.data());
            bad += memcmp(vg.data(), vc.data(), vg.size() * sizeof(sr_vertex)) != 0;
            std::vector<uint32_t> tg((size_t)(w - 1) * (h - 1) * 6), tc(tg.size());
            uint32_t ng = 0, nc = 0;
            CHECK(sr_mesh_faces(ctx, vg.data(), w, h, 2.0f, tg.data(), &ng));
            orc_mesh_faces(vc.data(), w, h, 2.0f, tc.data(), &nc);
            bad += ng != nc || memcmp(tg.data(), tc.data(), (size_t)ng * 12) != 0;
            printf("reconstruction chain: %u faces (oracle %u)\n", ng, nc);
        }
    }
    // error convention: h1 < h2 overruns the left ROI -> SR_E_SHAPE (cv::Exception in the reference)
    {
        sr_image ls{L.data(), w, h - 20, (size_t)w * 3};
        sr_search_params p{SR_VARIANT_RIGHT, 9, 0, D, 1.0, 0, 19.0, SR_COST_SSD_L2, 0, 0};
        std::vector<double> o((size_t)w * h);
        int rc = sr_block_search(ctx, &ls, &ri, &p, o.data(), (size_t)w * 8, nullptr);
        printf("h1<h2 -> rc %d (%s)\n", rc, sr_last_error(ctx));
        bad += rc != SR_E_SHAPE;
    }
    printf("kernel launches: %llu\n", (unsigned long long)sr_kernel_launches(ctx));
    sr_ctx_destroy(ctx);
    printf(bad ? "FAILED\n" : "OK\n");
    return bad ? 1 : 0;
}
