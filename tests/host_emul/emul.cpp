// TEST INFRASTRUCTURE ONLY — CPU build of the product's device functions (traverse.cuh, shade.cuh, scene_build.cpp)
// through the BLOKB_HD shim in blok_b200/csrc/hd.h. It exists so that the DDA's logic (not its performance) can be
// checked against the oracle in a container without a GPU, and so that world-size-2 sharding tests can run on CPU.
// Nothing in the product loads this library; the product has no CPU path.
#include <atomic>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "../../blok_b200/csrc/scene_build.h"
#include "../../blok_b200/csrc/post.cuh"
#include "../../blok_b200/csrc/shade.cuh"

using namespace blokb;

namespace {
struct Emu {
    HostWorld hw;
    DeviceWorld dw{};
    std::vector<blokb_material> mats;
};
template <class F>
void pfor(size_t n, int nt, F&& fn) {
    if (nt <= 1) { fn(0, n); return; }
    std::vector<std::thread> th; std::atomic<size_t> next{0};
    for (int t = 0; t < nt; t++) th.emplace_back([&] { for (;;) { size_t b = next.fetch_add(4096); if (b >= n) break; fn(b, std::min(n, b + 4096)); } });
    for (auto& x : th) x.join();
}
}  // namespace

extern "C" {
void* emu_create(const blokb_svo_node* nodes, size_t nNodes, const blokb_subchunk* subs, size_t nSubs, const blokb_material* mats, size_t nMat,
                 char* errOut, size_t errCap, int* rcOut) {
    Emu* e = new Emu;
    std::string err;
    int rc = build_host_world(nodes, nNodes, subs, nSubs, e->hw, err);
    if (rcOut) *rcOut = rc;
    if (rc) { if (errOut && errCap) { std::strncpy(errOut, err.c_str(), errCap - 1); errOut[errCap - 1] = 0; } delete e; return nullptr; }
    HostWorld& h = e->hw; DeviceWorld& d = e->dw;
    for (int a = 0; a < 3; a++) { d.origin[a] = h.origin[a]; d.dim[a] = h.dim[a]; }
    d.levels = h.levels; d.matWide = h.matWide;
    d.nodes = reinterpret_cast<const uint4*>(h.nodes.data()); d.blockFirstLeaf = h.blockFirstLeaf.data();
    d.voxelMat = h.matWide ? (const void*)h.mat32.data() : (const void*)h.mat16.data();
    e->mats.assign(mats, mats + nMat);
    if (e->mats.empty()) e->mats.push_back(blokb_material{{0.8f, 0.8f, 0.8f}, 127u << 16, {0, 0, 0}, 0.f});
    d.materials = reinterpret_cast<const float4*>(e->mats.data()); d.nMaterials = (uint32_t)e->mats.size();
    return e;
}
void emu_destroy(void* p) { delete (Emu*)p; }
void emu_stats(void* p, uint64_t* out /* bricks, blocks, voxels, compactBytes, levels, matWide */) {
    Emu* e = (Emu*)p;
    out[0] = e->hw.bricks.size(); out[1] = e->hw.blockFirstLeaf.size(); out[2] = e->hw.voxels; out[3] = e->hw.compactBytes(); out[4] = e->hw.levels; out[5] = e->hw.matWide;
}
// rays n x 8 -> blokb_hit n; returns total DDA steps
uint64_t emu_trace(void* p, const float* rays, size_t n, blokb_hit* out, int nThreads, int locate) {
    Emu* e = (Emu*)p;
    std::atomic<uint64_t> total{0};
    pfor(n, nThreads, [&](size_t b, size_t en) {
        uint64_t st = 0;
        for (size_t i = b; i < en; i++) {
            const float* r = rays + 8 * i;
            HitRec h; uint32_t steps = 0; uint4 stackMem[kMaxLevels]; const NodeStack stk{stackMem, 1};
            const bool hit = traverse<true>(e->dw, stk, r[0], r[1], r[2], r[3], r[4], r[5], r[6], r[7], h, steps, locate != 0);
            if (hit) out[i] = blokb_hit{h.node, h.material, h.t, h.facePrim};
            else out[i] = blokb_hit{0xFFFFFFFFu, 0xFFFFFFFFu, -1.0f, 0xFFFFFFFFu};
            st += steps;
        }
        total += st;
    });
    return total;
}
// Ray i of `second` traced right after ray i of `first` by the same "thread", starting from where the first walk left the node stack
// (WalkResume) — the way the kernels trace a sun shadow / bounce ray after the camera ray. out = the second rays' records; they must
// equal emu_trace(second) bit for bit. Returns how many second rays found a usable stack (started below the root).
uint64_t emu_trace_chain(void* p, const float* first, const float* second, size_t n, blokb_hit* out, int locateFirst) {
    Emu* e = (Emu*)p;
    uint64_t resumed = 0;
    for (size_t i = 0; i < n; i++) {
        uint4 stackMem[kMaxLevels]; const NodeStack stk{stackMem, 1};
        HitRec h; uint32_t steps = 0;
        WalkResume rs;
        const float* a = first + 8 * i; const float* b = second + 8 * i;
        traverse<false>(e->dw, stk, a[0], a[1], a[2], a[3], a[4], a[5], a[6], a[7], h, steps, locateFirst != 0, &rs);
        resumed += rs.level > 0;
        const bool hit = traverse<false>(e->dw, stk, b[0], b[1], b[2], b[3], b[4], b[5], b[6], b[7], h, steps, true, &rs);
        if (hit) out[i] = blokb_hit{h.node, h.material, h.t, h.facePrim};
        else out[i] = blokb_hit{0xFFFFFFFFu, 0xFFFFFFFFu, -1.0f, 0xFFFFFFFFu};
    }
    return resumed;
}
// Where the walk spends its decisions (profiling aid, tools/walk_profile.py): the loop of traverse() restated around the
// Walker's own methods, counting per child-cell size (index = sh / 2: 0 = voxels, 1 = 4^3 bricks, ...) the descents into a
// cell of that child size, how many of them needed plane tests (select) rather than m's bits, the steps taken at that size
// and the climbs that landed there. prof = 4 x (kMaxLevels + 1) counters; returns the number of hits.
uint64_t emu_walk_profile(void* p, const float* rays, size_t n, int locate, uint64_t* prof) {
    Emu* e = (Emu*)p;
    constexpr int K = kMaxLevels + 1;
    uint64_t hits = 0;
    for (size_t i = 0; i < n; i++) {
        const float* r = rays + 8 * i;
        uint4 stackMem[kMaxLevels]; const NodeStack stk{stackMem, 1};
        Walker wk;
        if (!wk.init(e->dw, r[0], r[1], r[2], r[3], r[4], r[5], r[6], r[7], locate != 0)) continue;
        for (;;) {
            bool hit = false;
            while (wk.occupied()) {
                if (wk.sh == 0) { hit = true; break; }
                const bool tests = wk.sh - 2 < wk.rs;
                wk.descend(e->dw, stk);
                prof[0 * K + wk.sh / 2]++;
                if (tests) prof[1 * K + wk.sh / 2]++;
            }
            if (hit) { hits++; break; }
            const int before = wk.sh;
            prof[2 * K + wk.sh / 2]++;
            if (!wk.step(e->dw, stk)) break;
            if (wk.sh != before) prof[3 * K + wk.sh / 2]++;
        }
    }
    return hits;
}
// The same loop run by 32 Walkers in lock step, the way a warp executes traverse(): rays[32 * w + l] is lane l of warp w. For each
// decision kind the counters hold {warp-level issues, lane-level executions}: utilisation = lanes / (32 * issues).
// out[0..1] descend, [2..3] select inside descend, [4..5] step, [6..7] widen inside step, [8..9] climb inside step, [10] warps.
// maxDescPasses <= 0: the product's loop shape (descend until no lane sits on an occupied cell, then step). > 0: a what-if shape in
// which at most that many descend passes precede each step pass and lanes still on an occupied cell skip the step.
void emu_warp_profile(void* p, const float* rays, size_t nWarps, int locate, uint64_t* out, int maxDescPasses) {
    Emu* e = (Emu*)p;
    for (size_t w = 0; w < nWarps; w++) {
        Walker wk[32]; uint4 stackMem[32][kMaxLevels]; bool live[32];
        for (int l = 0; l < 32; l++) {
            const float* r = rays + 8 * (32 * w + l);
            live[l] = wk[l].init(e->dw, r[0], r[1], r[2], r[3], r[4], r[5], r[6], r[7], locate != 0);
        }
        out[10]++;
        for (;;) {
            for (int pass = 0; maxDescPasses <= 0 || pass < maxDescPasses; pass++) {  // while (occupied) descend: one issue per pass in which any lane descends
                uint32_t nDesc = 0, nSel = 0;
                for (int l = 0; l < 32; l++) {
                    if (!live[l] || !wk[l].occupied()) continue;
                    if (wk[l].sh == 0) { live[l] = false; continue; }  // hit: leaves the loop, waits for the warp
                    nSel += wk[l].sh - 2 < wk[l].rs;
                    wk[l].descend(e->dw, NodeStack{stackMem[l], 1});
                    nDesc++;
                }
                if (!nDesc) break;
                out[0]++; out[1] += nDesc;
                if (nSel) { out[2]++; out[3] += nSel; }
            }
            uint32_t nStep = 0, nWiden = 0, nClimb = 0;
            uint32_t nWait = 0;
            for (int l = 0; l < 32; l++) {
                if (!live[l]) continue;
                if (wk[l].occupied()) {  // only with a bounded descend loop (a what-if shape): still descending, or a hit found late
                    if (wk[l].sh == 0) live[l] = false; else nWait++;
                    continue;
                }
                nStep++;
                nWiden += wk[l].sh > wk[l].rs;
                const int before = wk[l].sh;
                if (!wk[l].step(e->dw, NodeStack{stackMem[l], 1})) { live[l] = false; continue; }
                nClimb += wk[l].sh != before;
            }
            if (!nStep && !nWait) break;
            if (!nStep) continue;
            out[4]++; out[5] += nStep;
            if (nWiden) { out[6]++; out[7] += nWiden; }
            if (nClimb) { out[8]++; out[9] += nClimb; }
        }
    }
}
// What-if schedule for the same 32 lock-step Walkers: ONE pass per iteration, chosen by vote — a descend pass when the lanes sitting on
// an occupied cell number at least num/den of the lanes wanting a step, else a step pass; the minority waits. out as emu_warp_profile
// plus out[11] = iterations (each costs the vote's own instructions).
void emu_warp_vote(void* p, const float* rays, size_t nWarps, int locate, int num, int den, uint64_t* out) {
    Emu* e = (Emu*)p;
    for (size_t w = 0; w < nWarps; w++) {
        Walker wk[32]; uint4 stackMem[32][kMaxLevels]; bool live[32];
        for (int l = 0; l < 32; l++) {
            const float* r = rays + 8 * (32 * w + l);
            live[l] = wk[l].init(e->dw, r[0], r[1], r[2], r[3], r[4], r[5], r[6], r[7], locate != 0);
        }
        out[10]++;
        for (;;) {
            uint32_t nd = 0, ns = 0;
            for (int l = 0; l < 32; l++) {
                if (!live[l]) continue;
                if (wk[l].occupied()) { if (wk[l].sh == 0) live[l] = false; else nd++; }
                else ns++;
            }
            if (!nd && !ns) break;
            out[11]++;
            if (nd && (ns == 0 || uint64_t(nd) * den >= uint64_t(ns) * num)) {
                uint32_t nSel = 0;
                for (int l = 0; l < 32; l++) {
                    if (!live[l] || !wk[l].occupied()) continue;
                    nSel += wk[l].sh - 2 < wk[l].rs;
                    wk[l].descend(e->dw, NodeStack{stackMem[l], 1});
                }
                out[0]++; out[1] += nd;
                if (nSel) { out[2]++; out[3] += nSel; }
            } else {
                uint32_t nWiden = 0, nClimb = 0;
                for (int l = 0; l < 32; l++) {
                    if (!live[l] || wk[l].occupied()) continue;
                    nWiden += wk[l].sh > wk[l].rs;
                    const int before = wk[l].sh;
                    if (!wk[l].step(e->dw, NodeStack{stackMem[l], 1})) { live[l] = false; continue; }
                    nClimb += wk[l].sh != before;
                }
                out[4]++; out[5] += ns;
                if (nWiden) { out[6]++; out[7] += nWiden; }
                if (nClimb) { out[8]++; out[9] += nClimb; }
            }
        }
    }
}
// per-ray DDA step counts (profiling aid for warp-divergence estimates)
void emu_trace_steps(void* p, const float* rays, size_t n, uint32_t* stepsOut, int nThreads) {
    Emu* e = (Emu*)p;
    pfor(n, nThreads, [&](size_t b, size_t en) {
        for (size_t i = b; i < en; i++) {
            const float* r = rays + 8 * i;
            HitRec h; uint32_t steps = 0; uint4 stackMem[kMaxLevels]; const NodeStack stk{stackMem, 1};
            traverse<true>(e->dw, stk, r[0], r[1], r[2], r[3], r[4], r[5], r[6], r[7], h, steps);
            stepsOut[i] = steps;
        }
    });
}
// radiance sum of samples [sBegin, sEnd) per pixel (k_path without the accumulate), out = w*h*4 floats (sum.rgb, count)
void emu_path(void* p, const blokb_frame_ubo* f, uint32_t sBegin, uint32_t sEnd, float* out, int nThreads) {
    Emu* e = (Emu*)p;
    FrameDev F;
    std::memcpy(F.invView, f->invView, 64); std::memcpy(F.invProj, f->invProj, 64); std::memcpy(F.prevViewProj, f->prevViewProj, 64);
    std::memcpy(F.camPos, f->camPos, 12);
    F.frameCount = f->frameCount; F.sampleCount = f->sampleCount; F.width = f->screenWidth; F.height = f->screenHeight;
    const size_t n = size_t(F.width) * F.height;
    pfor(n, nThreads, [&](size_t b, size_t en) {
        for (size_t i = b; i < en; i++) {
            uint32_t px = uint32_t(i % F.width), py = uint32_t(i / F.width), rays = 0, steps = 0;
            F3 acc = f3(0.f); uint4 stackMem[kMaxLevels]; const NodeStack stk{stackMem, 1};
            for (uint32_t s = sBegin; s < sEnd; s++) acc = acc + path_sample<false>(e->dw, stk, F, px, py, s, nullptr, rays, steps);
            out[4 * i] = acc.x; out[4 * i + 1] = acc.y; out[4 * i + 2] = acc.z; out[4 * i + 3] = float(sEnd - sBegin);
        }
    });
}

// ---- denoiser / post-process pixel functions (post.cuh), looped over the image on the host --------------------
void emu_post_frame(const blokb_frame_ubo* u, const blokb_post_settings* s, post::PostFrame* out) { *out = post::make_post_frame(*u, *s); }
size_t emu_post_frame_size() { return sizeof(post::PostFrame); }
void emu_post_temporal_variance(const post::PostFrame* P, const post::TemporalIO* io, int nThreads) {
    pfor(size_t(P->width) * P->height, nThreads, [&](size_t b, size_t e) {
        for (size_t i = b; i < e; i++) post::temporal_variance_pixel(*P, *io, int(i % P->width), int(i / P->width));
    });
}
void emu_post_atrous(const post::PostFrame* P, int step, const float* in, const float* var, const float* wp, const uint16_t* nr, float* out, int nThreads) {
    pfor(size_t(P->width) * P->height, nThreads, [&](size_t b, size_t e) {
        for (size_t i = b; i < e; i++) post::atrous_pixel(*P, step, in, var, wp, nr, out, int(i % P->width), int(i / P->width));
    });
}
void emu_post_taa_tonemap(const post::PostFrame* P, int taa, int tonemap, const float* cur, const float* prevHist, const uint16_t* motion, float* outHist,
                          float* outColor, uint32_t* outLdr, int nThreads) {
    pfor(size_t(P->width) * P->height, nThreads, [&](size_t b, size_t e) {
        for (size_t i = b; i < e; i++) {
            const int x = int(i % P->width), y = int(i / P->width);
            if (taa && tonemap) post::taa_tonemap_pixel<true, true>(*P, cur, prevHist, motion, outHist, outColor, outLdr, x, y);
            else if (taa) post::taa_tonemap_pixel<true, false>(*P, cur, prevHist, motion, outHist, outColor, outLdr, x, y);
            else post::taa_tonemap_pixel<false, true>(*P, cur, prevHist, motion, outHist, outColor, outLdr, x, y);
        }
    });
}
void emu_post_sharpen(const post::PostFrame* P, const uint8_t* in, uint32_t* out, int nThreads) {
    float lut255[256];
    for (int i = 0; i < 256; i++) lut255[i] = float(i) / 255.0f;
    pfor(size_t(P->width) * P->height, nThreads, [&](size_t b, size_t e) {
        for (size_t i = b; i < e; i++) post::sharpen_pixel(*P, lut255, in, out, int(i % P->width), int(i / P->width));
    });
}
}
