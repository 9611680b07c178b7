// session_bench.cpp — config C5 in its real-time form: S desktop sessions per GPU on G GPUs, every
// session delivering a 1080p BGRx frame at 60 Hz, one worker thread per GPU converting each tick's
// frames with one b200vc_convert_batch call. Reports, per GPU and overall, the frames converted and
// the latency of a tick (all of that GPU's sessions converted and back in host memory).
//
// Sessions share nothing (the reference runs one session per container per GPU, /root/reference
// xgl.yml:10,120; README.md:182), so GPUs never talk: one libb200vc context per GPU, no collective.
// Plain C++ over the C ABI; no Python, no torch.
//
//   session_bench [--gpus G] [--device D] [--sessions S] [--seconds T] [--hz F] [--damage]
//                 [--in WxH] [--out WxH] [--nv12] [--matrix 0|1|2] [--full] [--wait-file PATH]
// --hz 0 runs the ticks back to back (end-to-end throughput with no interpreter in the loop);
// --device D puts a one-GPU run on device D (one process per GPU, the way bench.py's ranks use it);
// --wait-file makes every worker start its first tick only once PATH exists (ranks start together).
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include <unistd.h>

#include "b200vc.h"

using clk = std::chrono::steady_clock;

struct GpuResult {
    int device = 0;
    int rc = 0;
    std::string err;
    uint64_t frames = 0, ticks = 0, late_ticks = 0;
    double run_s = 0;                      // first tick to last tick (set-up excluded)
    std::vector<float> tick_ms;
};

static uint64_t splitmix64(uint64_t &s)
{
    uint64_t z = (s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

struct Geometry { int W = 1920, H = 1080, OW = 1920, OH = 1080, fmt = B200VC_FMT_I420, matrix = -1, range = -1; };
static std::string g_wait_file;

static void worker(int device, int sessions, double seconds, double hz, bool damage, Geometry geo, GpuResult *res)
{
    res->device = device;
    const int W = geo.W, H = geo.H;
    b200vc_config cfg;
    memset(&cfg, 0, sizeof cfg);
    cfg.device = device;
    cfg.in_w = W;  cfg.in_h = H;
    cfg.out_w = geo.OW;  cfg.out_h = geo.OH;
    cfg.out_format = geo.fmt;
    b200vc_default_colorimetry(geo.OH, &cfg.matrix, &cfg.range, &cfg.chroma_site);
    if (geo.matrix >= 0) cfg.matrix = geo.matrix;
    if (geo.range >= 0) cfg.range = geo.range;
    cfg.max_sessions = sessions;
    cfg.n_inflight = sessions < 64 ? sessions : 64;   // (the context keeps one slot per session anyway)
    cfg.flags = damage ? B200VC_FLAG_DAMAGE_MAP : 0;
    b200vc_ctx *ctx = nullptr;
    if ((res->rc = b200vc_create(&cfg, &ctx)) != B200VC_OK) {
        res->err = b200vc_last_error(nullptr);
        return;
    }
    size_t off[3], pit[3];
    const size_t out_bytes = b200vc_default_layout(cfg.out_format, geo.OW, geo.OH, off, pit);
    const size_t in_bytes = (size_t)W * H * 4, mb = (size_t)((W + 15) / 16) * ((H + 15) / 16);
    // two input frames per session (a desktop that changes a little every tick), pinned
    std::vector<uint8_t *> in(sessions * 2), out(sessions), dmg(sessions);
    for (int s = 0; s < sessions; s++) {
        for (int k = 0; k < 2; k++) {
            void *p = nullptr;
            if (b200vc_host_alloc(in_bytes, &p) != B200VC_OK) { res->rc = B200VC_ERR_NOMEM; res->err = "pinned allocation failed"; return; }
            in[s * 2 + k] = (uint8_t *)p;
        }
        void *p = nullptr;
        if (b200vc_host_alloc(out_bytes + mb, &p) != B200VC_OK) { res->rc = B200VC_ERR_NOMEM; res->err = "pinned allocation failed"; return; }
        out[s] = (uint8_t *)p;
        dmg[s] = out[s] + out_bytes;
        uint64_t st = 0x1234 + 977 * s + 131071 * device;
        for (size_t i = 0; i < in_bytes; i += 8) { uint64_t v = splitmix64(st); memcpy(in[s * 2] + i, &v, 8); }
        memcpy(in[s * 2 + 1], in[s * 2], in_bytes);
        for (int y = 100; y < 124; y++) memset(in[s * 2 + 1] + ((size_t)y * W + 200 + 7 * s) * 4, 0xff, 16 * 4);   // a "cursor"
    }
    std::vector<b200vc_frame> frames(sessions);
    // two untimed ticks: kernel images loaded, staging and flags touched, before the clock starts (a session has
    // been running for a while when its 60 Hz budget matters)
    for (int warm = 0; warm < 2; warm++) {
        for (int s = 0; s < sessions; s++) {
            b200vc_frame &f = frames[s];
            memset(&f, 0, sizeof f);
            f.bgrx = in[s * 2 + (warm & 1)];
            f.in_pitch = (size_t)W * 4;
            for (int p = 0; p < 3; p++) { f.planes[p] = out[s] + off[p]; f.pitches[p] = pit[p]; }
            f.damage = damage ? dmg[s] : nullptr;
            f.session = s;
        }
        if ((res->rc = b200vc_convert_batch(ctx, frames.data(), sessions)) != B200VC_OK) {
            res->err = b200vc_last_error(ctx);
            return;
        }
    }
    if (!g_wait_file.empty())
        while (access(g_wait_file.c_str(), F_OK) != 0) std::this_thread::sleep_for(std::chrono::microseconds(200));
    const bool paced = hz > 0;
    const auto period = std::chrono::duration_cast<clk::duration>(std::chrono::duration<double>(paced ? 1.0 / hz : 0.0));
    const auto t_end = clk::now() + std::chrono::duration_cast<clk::duration>(std::chrono::duration<double>(seconds));
    auto next = clk::now();
    const auto t_run = next;
    for (uint64_t tick = 0; clk::now() < t_end; tick++) {
        if (paced) std::this_thread::sleep_until(next);
        const auto t0 = clk::now();
        for (int s = 0; s < sessions; s++) {
            b200vc_frame &f = frames[s];
            memset(&f, 0, sizeof f);
            f.bgrx = in[s * 2 + (tick & 1)];
            f.in_pitch = (size_t)W * 4;
            for (int p = 0; p < 3; p++) { f.planes[p] = out[s] + off[p]; f.pitches[p] = pit[p]; }
            f.damage = damage ? dmg[s] : nullptr;
            f.session = s;
        }
        if ((res->rc = b200vc_convert_batch(ctx, frames.data(), sessions)) != B200VC_OK) {
            res->err = b200vc_last_error(ctx);
            break;
        }
        const float ms = std::chrono::duration<float, std::milli>(clk::now() - t0).count();
        res->tick_ms.push_back(ms);
        res->frames += sessions;
        res->ticks += 1;
        next += period;
        if (paced && clk::now() > next) { res->late_ticks += 1; next = clk::now(); }   // missed the period: resynchronise
    }
    res->run_s = std::chrono::duration<double>(clk::now() - t_run).count();
    b200vc_destroy(ctx);
    for (uint8_t *p : in) b200vc_host_free(p);
    for (uint8_t *p : out) b200vc_host_free(p);
}

static float pct(std::vector<float> v, double q)
{
    if (v.empty()) return 0;
    std::sort(v.begin(), v.end());
    return v[std::min(v.size() - 1, (size_t)(q * v.size()))];
}

int main(int argc, char **argv)
{
    int gpus = 1, sessions = 8, device0 = 0;
    double seconds = 5, hz = 60;
    bool damage = false;
    Geometry geo;
    for (int i = 1; i < argc; i++) {
        if (!strcmp(argv[i], "--gpus") && i + 1 < argc) gpus = atoi(argv[++i]);
        else if (!strcmp(argv[i], "--device") && i + 1 < argc) device0 = atoi(argv[++i]);
        else if (!strcmp(argv[i], "--in") && i + 1 < argc) { if (sscanf(argv[++i], "%dx%d", &geo.W, &geo.H) != 2) return 2; geo.OW = geo.W; geo.OH = geo.H; }
        else if (!strcmp(argv[i], "--out") && i + 1 < argc) { if (sscanf(argv[++i], "%dx%d", &geo.OW, &geo.OH) != 2) return 2; }
        else if (!strcmp(argv[i], "--nv12")) geo.fmt = B200VC_FMT_NV12;
        else if (!strcmp(argv[i], "--matrix") && i + 1 < argc) geo.matrix = atoi(argv[++i]);
        else if (!strcmp(argv[i], "--full")) geo.range = B200VC_RANGE_FULL;
        else if (!strcmp(argv[i], "--wait-file") && i + 1 < argc) g_wait_file = argv[++i];
        else if (!strcmp(argv[i], "--sessions") && i + 1 < argc) sessions = atoi(argv[++i]);
        else if (!strcmp(argv[i], "--seconds") && i + 1 < argc) seconds = atof(argv[++i]);
        else if (!strcmp(argv[i], "--hz") && i + 1 < argc) hz = atof(argv[++i]);
        else if (!strcmp(argv[i], "--damage")) damage = true;
        else { fprintf(stderr, "usage: %s [--gpus G] [--device D] [--sessions S] [--seconds T] [--hz F] [--damage] [--in WxH] [--out WxH] [--nv12] [--matrix M] [--full] [--wait-file PATH]\n", argv[0]); return 2; }
    }
    std::vector<GpuResult> res(gpus);
    std::vector<std::thread> th;
    const auto t0 = clk::now();
    for (int g = 0; g < gpus; g++) th.emplace_back(worker, device0 + g, sessions, seconds, hz, damage, geo, &res[g]);
    for (auto &t : th) t.join();
    const double wall = std::chrono::duration<double>(clk::now() - t0).count();
    uint64_t frames = 0, late = 0, ticks = 0;
    double run_s = 0;
    std::vector<float> all;
    int rc = 0;
    for (const GpuResult &r : res) {
        run_s = std::max(run_s, r.run_s);
        if (r.rc) { fprintf(stderr, "GPU %d: status %d: %s\n", r.device, r.rc, r.err.c_str()); rc = 1; }
        frames += r.frames;  late += r.late_ticks;  ticks += r.ticks;
        all.insert(all.end(), r.tick_ms.begin(), r.tick_ms.end());
    }
    printf("{\"config\": \"%d sessions x %d GPUs, %dx%d BGRx->%s %dx%d, %s%s\", \"sessions\": %d, \"frames\": %llu, "
           "\"wall_s\": %.2f, \"aggregate_fps\": %.1f, \"target_fps\": %.1f, \"ticks\": %llu, \"late_ticks\": %llu, "
           "\"tick_ms_p50\": %.3f, \"tick_ms_p99\": %.3f, \"tick_ms_max\": %.3f, \"tick_budget_ms\": %.3f, \"per_gpu_p99_ms\": [",
           sessions, gpus, geo.W, geo.H, geo.fmt == B200VC_FMT_NV12 ? "NV12" : "I420", geo.OW, geo.OH,
           hz > 0 ? (std::to_string((int)hz) + " Hz").c_str() : "back to back", damage ? ", damage map" : "",
           sessions * gpus, (unsigned long long)frames, wall, run_s > 0 ? frames / run_s : 0.0,
           sessions * gpus * hz, (unsigned long long)ticks, (unsigned long long)late, pct(all, 0.5), pct(all, 0.99), pct(all, 0.9999),
           hz > 0 ? 1000.0 / hz : 0.0);
    for (int g = 0; g < gpus; g++) printf("%s%.3f", g ? ", " : "", pct(res[g].tick_ms, 0.99));
    printf("]}\n");
    return rc;
}
