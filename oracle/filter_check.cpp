// filter_check.cpp — TEST INFRASTRUCTURE.  Brute-force check that the sweep's conservative candidate filter
// (rust-ray-tracing_b200/csrc/rt_filter.h, compiled here for the host with the same explicit-fma arithmetic) never
// rejects a (ray, sphere) pair whose reference-order float discriminant (rt_oracle.hpp sphere_discriminant<float>,
// src/objects.rs:252-257) is >= 0.  Pairs are adversarial: directions aimed at the silhouette of the sphere with
// relative perturbations from 1e-9 to 1e-1, origins at the camera, on other spheres' surfaces (bounce rays) and far
// out on the ground sphere, direction lengths 0.1..2 (the reference never normalises).
//
// usage: filter_check <scenario 0..2> <pairs> <seed> <threads>   -> one JSON line on stdout
#include "rt_oracle.hpp"
#include "../rust-ray-tracing_b200/csrc/rt_filter.h"

#include <atomic>
#include <cinttypes>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <thread>
#include <vector>

using namespace rto;

struct Sph { float cx, cy, cz, r; };

struct Tally {
    uint64_t pairs = 0, ref_pos = 0, flagged = 0, false_neg = 0;
    uint64_t slab_flagged[3] = {0, 0, 0}, slab_false_neg = 0;   // stage 1 (slab test), per dropped axis
    double min_slack = 1e300;   // min over reference-positive pairs of q / (ray margin + sphere margin)
};

static std::vector<Sph> make_scene(int scenario, std::mt19937_64& g) {
    std::uniform_real_distribution<double> U(0.0, 1.0);
    std::vector<Sph> s;
    if (scenario == 0) {   // the random-spheres layout
        s.push_back({0.f, -1000.f, 0.f, 1000.f});
        for (int a = -11; a < 11; a++)
            for (int b = -11; b < 11; b++) s.push_back({(float)(a + 0.9 * U(g)), 0.2f, (float)(b + 0.9 * U(g)), 0.2f});
        s.push_back({0.f, 1.f, 0.f, 1.f}); s.push_back({-4.f, 1.f, 0.f, 1.f}); s.push_back({4.f, 1.f, 0.f, 1.f});
    } else if (scenario == 1) {   // 10k grid (S3) extents
        s.push_back({0.f, -1000.f, 0.f, 1000.f});
        for (int i = 0; i < 600; i++) s.push_back({(float)(-50 + 100 * U(g)), 0.2f, (float)(-50 + 100 * U(g)), 0.2f});
    } else {   // wild: radii 1e-3..1e3, centres up to 1e4 from the origin
        for (int i = 0; i < 500; i++) {
            double sc = std::pow(10.0, -1 + 5 * U(g));
            double r = std::pow(10.0, -3 + 6 * U(g));
            s.push_back({(float)(sc * (2 * U(g) - 1)), (float)(sc * (2 * U(g) - 1)), (float)(sc * (2 * U(g) - 1)), (float)r});
        }
    }
    return s;
}

static void unit_random(std::mt19937_64& g, double v[3]) {
    std::normal_distribution<double> N(0.0, 1.0);
    double l;
    do { v[0] = N(g); v[1] = N(g); v[2] = N(g); l = std::sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]); } while (l < 1e-6);
    v[0] /= l; v[1] /= l; v[2] /= l;
}

static void worker(int scenario, uint64_t pairs, uint64_t seed, Tally* out) {
    std::mt19937_64 g(seed);
    std::uniform_real_distribution<double> U(0.0, 1.0);
    Tally t;
    std::vector<Sph> scene = make_scene(scenario, g);
    std::vector<float> k(scene.size()), rp(scene.size());
    for (size_t i = 0; i < scene.size(); i++) {
        k[i] = rtb::filter_sphere_k(scene[i].cx, scene[i].cy, scene[i].cz, scene[i].r * scene[i].r);
        rp[i] = rtb::slab_sphere_r(scene[i].cx, scene[i].cy, scene[i].cz, scene[i].r * scene[i].r);
    }
    const int n = (int)scene.size();
    const float rp_min = rtb::slab_apply_floor(rp.data(), n, 1);   // as rtb200_upload_scene does
    while (t.pairs < pairs) {
        // --- origin
        double o[3];
        double pick = U(g);
        if (pick < 0.25) { o[0] = 16; o[1] = 2; o[2] = 18.5; }
        else {
            const Sph& b = (pick < 0.45) ? scene[0] : scene[(size_t)(U(g) * n) % n];   // bounce off the big one / any sphere
            double nn[3]; unit_random(g, nn);
            if (&b == &scene[0] && scenario < 2) {   // near the top of the ground sphere, out to ~600 units
                double rad = std::pow(10.0, -1 + 3.8 * U(g)), ang = 6.283185307179586 * U(g);
                double x = rad * std::cos(ang), z = rad * std::sin(ang);
                if (rad < 999.0) { nn[0] = x / 1000.0; nn[2] = z / 1000.0; nn[1] = std::sqrt(1 - nn[0] * nn[0] - nn[2] * nn[2]); }
            }
            for (int q = 0; q < 3; q++) o[q] = (&b.cx)[q] + nn[q] * b.r;
        }
        float of[3] = {(float)o[0], (float)o[1], (float)o[2]};
        // --- target sphere and a near-silhouette direction
        int ti = (int)((size_t)(U(g) * n) % n);
        const Sph& s = scene[ti];
        double oc[3] = {s.cx - (double)of[0], s.cy - (double)of[1], s.cz - (double)of[2]};
        double L = std::sqrt(oc[0] * oc[0] + oc[1] * oc[1] + oc[2] * oc[2]);
        double d[3];
        if (L > s.r * (1 + 1e-9) && U(g) < 0.9) {
            double ax[3]; unit_random(g, ax);
            double dp = (ax[0] * oc[0] + ax[1] * oc[1] + ax[2] * oc[2]) / (L * L);
            for (int q = 0; q < 3; q++) ax[q] -= dp * oc[q];
            double al = std::sqrt(ax[0] * ax[0] + ax[1] * ax[1] + ax[2] * ax[2]);
            if (al < 1e-9) continue;
            double eps = std::pow(10.0, -9 + 8 * U(g)) * (U(g) < 0.5 ? -1 : 1);
            double theta = std::asin(s.r / L) * (1 + eps);
            for (int q = 0; q < 3; q++) d[q] = std::cos(theta) * oc[q] / L + std::sin(theta) * ax[q] / al;
        } else {
            unit_random(g, d);
        }
        double scale = 0.1 + 1.9 * U(g);
        Ray<float> ray{{of[0], of[1], of[2]}, {(float)(d[0] * scale), (float)(d[1] * scale), (float)(d[2] * scale)}};
        float a = packed_length_squared(ray.d);
        rtb::FilterRay fr = rtb::make_filter_ray(ray.o.x, ray.o.y, ray.o.z, ray.d.x, ray.d.y, ray.d.z, a);
        rtb::SlabRay sr[3];
        for (int up = 0; up < 3; up++) sr[up] = rtb::make_slab_ray(ray.o.x, ray.o.y, ray.o.z, ray.d.x, ray.d.y, ray.d.z, a, up, rp_min);
        double o2 = (double)of[0] * of[0] + (double)of[1] * of[1] + (double)of[2] * of[2];
        // --- the aimed-at sphere plus a handful of others per ray
        for (int rep = 0; rep < 8 && t.pairs < pairs; rep++) {
            int si = rep == 0 ? ti : (int)((size_t)(U(g) * n) % n);
            const Sph& sp = scene[si];
            Sphere<float> os{{sp.cx, sp.cy, sp.cz}, sp.r, 0};
            float hb, aa;
            float disc = sphere_discriminant(os, ray, hb, aa);
            float q = rtb::filter_value(fr, sp.cx, sp.cy, sp.cz, k[si]);
            bool ref = disc >= 0.0f;
            bool flag = !std::signbit(q);
            t.pairs++;
            t.ref_pos += ref;
            t.flagged += flag;
            if (ref && !flag) t.false_neg++;
            for (int up = 0; up < 3; up++) {
                const float cc[3] = {sp.cx, sp.cy, sp.cz};
                const int ia = (up + 1) % 3, ib = (up + 2) % 3;
                bool sflag = !std::signbit(rtb::slab_value(sr[up], cc[ia], cc[ib], rp[si]));
                t.slab_flagged[up] += sflag;
                if (ref && !sflag) t.slab_false_neg++;
            }
            if (ref) {
                double b = std::sqrt((double)sp.cx * sp.cx + (double)sp.cy * sp.cy + (double)sp.cz * sp.cz) + sp.r;
                double margin = (double)RTB_FILTER_MARGIN * (o2 + b * b);
                if (margin > 0) t.min_slack = std::min(t.min_slack, (double)q / margin);
            }
        }
    }
    *out = t;
}

int main(int argc, char** argv) {
    int scenario = argc > 1 ? std::atoi(argv[1]) : 0;
    uint64_t pairs = argc > 2 ? std::strtoull(argv[2], nullptr, 10) : 10000000ull;
    uint64_t seed = argc > 3 ? std::strtoull(argv[3], nullptr, 10) : 1;
    int threads = argc > 4 ? std::atoi(argv[4]) : 1;
    if (threads < 1) threads = 1;
    std::vector<Tally> tl(threads);
    std::vector<std::thread> th;
    for (int i = 0; i < threads; i++) th.emplace_back(worker, scenario, pairs / threads, seed * 7919 + i, &tl[i]);
    for (auto& x : th) x.join();
    Tally t;
    for (auto& x : tl) {
        t.pairs += x.pairs; t.ref_pos += x.ref_pos; t.flagged += x.flagged; t.false_neg += x.false_neg;
        for (int up = 0; up < 3; up++) t.slab_flagged[up] += x.slab_flagged[up];
        t.slab_false_neg += x.slab_false_neg;
        t.min_slack = std::min(t.min_slack, x.min_slack);
    }
    std::printf("{\"scenario\": %d, \"pairs\": %" PRIu64 ", \"ref_positive\": %" PRIu64 ", \"flagged\": %" PRIu64
                ", \"false_negatives\": %" PRIu64 ", \"min_slack\": %.6g, \"slab_flagged\": [%" PRIu64 ", %" PRIu64 ", %" PRIu64
                "], \"slab_false_negatives\": %" PRIu64 "}\n",
                scenario, t.pairs, t.ref_pos, t.flagged, t.false_neg, t.min_slack, t.slab_flagged[0], t.slab_flagged[1], t.slab_flagged[2],
                t.slab_false_neg);
    return (t.false_neg == 0 && t.slab_false_neg == 0) ? 0 : 1;
}
