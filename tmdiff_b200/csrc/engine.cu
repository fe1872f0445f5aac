// Engine behind the C ABI (include/tmdiff_b200.h): weights, workspace, the WavBEST data-flow as
// a list of fused kernel launches, the step-invariant condition cache and the sampler loop.
//
// Data-flow restated from GeneralModel/Hyper_unet_general.py:600-636 (WavBEST.forward) and the
// block bodies :158-273, :334-414; see DESIGN.md for the fusion map.
#include <cuda_runtime.h>

#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <functional>
#include <map>
#include <string>
#include <vector>

#include "../../include/tmdiff_b200.h"
#include "kernels.h"

using namespace tmd;

static thread_local char g_err[1024] = "";
static int fail(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return 1;
}
#define CK(expr)                                                                                     \
    do {                                                                                             \
        cudaError_t _e = (expr);                                                                     \
        if (_e != cudaSuccess) return fail("%s:%d %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
    } while (0)

namespace {

struct View {  // channel slice of a channels-last activation tensor
    void* p = nullptr;
    int C = 0, ld = 0;
    bool f16 = false;   // bf16 mode: an add-only tensor kept as IEEE half (ConvEpilogue::raw_f16 / add0_f16)
};
struct ActSpec {
    View dst;
    const float* tb = nullptr;
    const float* sc = nullptr;
};
struct OutSpec {
    View raw;
    ActSpec act[2];
    View pool;   // 2x2 average of the result at the next level (consumed by a fused Conv_2, see ConvOp::src2)
    float* proj = nullptr;   // act[0] is not stored but projected onto one band plane with the engine's 32 -> 1 weights (ConvEpilogue::proj_out)
};
struct ConvW {
    float* w_f32 = nullptr;
    __nv_bfloat16* w_bf16 = nullptr;
    float* bias = nullptr;
    int Co = 0, Ci = 0, taps = 0;
};
struct HostTensor {
    std::vector<float> data;
    std::vector<int64_t> shape;
};

struct ProfSpan {
    cudaEvent_t a, b;
    int kind;
    double flops, bytes, flops_exec;
    const char* label;
};
using LaunchFn = std::function<cudaError_t(cudaStream_t)>;
struct Launch {       // one kernel launch of the plan + what it costs algorithmically (bench.py's roofline)
    LaunchFn fn;
    int kind = TMD_K_ELEMENTWISE;
    double flops = 0.0, bytes = 0.0;
    double flops_exec = -1.0;   // FLOPs the kernel issues (< flops where zero-padding work is skipped); -1: same as flops
    bool is_kernel = true;      // false: a memset / copy node (not counted by tmd_launch_count)
    std::string label;
    cudaError_t operator()(cudaStream_t s) const { return fn(s); }
};

}  // namespace

constexpr int kMaxPlanEvals = 256;
constexpr size_t kRedzoneBytes = 4096;   // evaluations per tmd_sample plan whose time tables are precomputed in one launch

struct tmd_engine {
    tmd_config cfg{};
    int dtype = 0;
    size_t esize = 4;
    int device = 0;
    std::map<std::string, HostTensor> host;
    std::map<std::string, std::vector<int64_t>> inv_shapes;
    std::map<std::string, ConvW> convs;
    std::map<std::string, int> time_off, text_off;
    int time_rows = 0, text_rows = 0;
    bool finalized = false, text_dirty = true;
    std::vector<void*> allocs;      // lives as long as the handle
    std::vector<void*> w_allocs;    // repacked weights and tables: released by the next tmd_finalize_weights
    bool weight_scope = false;

    // conditioning
    float *d_freqs = nullptr, *d_e_w1 = nullptr, *d_e_b1 = nullptr, *d_e_w2 = nullptr, *d_e_b2 = nullptr;
    float *d_p_w[3] = {}, *d_p_b[3] = {};
    float *d_Wt = nullptr, *d_bt = nullptr, *d_Ws = nullptr, *d_bs = nullptr;
    float *d_text = nullptr, *d_p_tmp0 = nullptr, *d_p_tmp1 = nullptr, *d_prompt = nullptr, *d_sc_sensor = nullptr;
    float *d_tb = nullptr, *d_sc = nullptr, *d_tin = nullptr, *d_embed = nullptr;
    int* d_sensor = nullptr;
    float *d_stem_w[2] = {}, *d_stem_b[2] = {}, *d_proj_w = nullptr;
    bool have_text[TMD_NUM_SENSORS] = {};
    std::vector<float> freqs_host;

    // bound geometry + launch lists
    int B = 0, N = 0, H = 0, W = 0;
    char* arena = nullptr;
    size_t arena_size = 0, arena_used = 0;
    std::map<std::string, void*> bufs;
    // TMD_ARENA_REDZONE=1 (debug): a canary-filled red zone behind every arena buffer, checked by tmd_debug_check_redzones
    // (compute-sanitizer is not available on every pool; this is the library's own bounds check of its workspace)
    std::vector<std::pair<std::string, char*>> redzones;
    bool redzone_on = false;
    std::vector<Launch> pre_ops, step_ops;
    std::vector<UmmaPlan*> plans;
    const float* cur_x = nullptr;
    const float* cur_pan = nullptr;
    const float* cur_ms = nullptr;
    float* cur_out = nullptr;
    const float* cur_t_dev = nullptr;
    float cur_t_scalar = 0.f;
    float* slots[TMD_NUM_SLOTS] = {};
    float* d_out = nullptr;
    float* d_xdup = nullptr;            // [x | x] for classifier-free-guidance plans (x evaluated against both condition halves)
    void* d_th_ws = nullptr;            // radix-select scratch of the dynamic threshold (solver.cu)
    // tmd_sample: the time tables of every evaluation of a plan are computed by ONE launch up front
    // (t is a host scalar per evaluation and shared by the batch); each evaluation then broadcasts its row.
    float *d_tb_all = nullptr, *d_tin_all = nullptr;
    const float* tb_bcast_src = nullptr;
    // tmd_sample: the launch sequence of a plan is captured once into a CUDA graph (on a private stream) and replayed,
    // which removes the per-launch gaps of ~1000 dependent kernels.  Keyed by the plan's contents; dropped on re-plan.
    struct SampleGraph { uint64_t key; cudaGraphExec_t exec; int64_t launches; };
    std::vector<SampleGraph> graphs;
    cudaStream_t cap_stream = nullptr;
    void drop_graphs() {
        for (auto& g : graphs) cudaGraphExecDestroy(g.exec);
        graphs.clear();
    }
    int64_t launches = 0;
    bool profiling = false;
    std::vector<ProfSpan> prof_events;
    size_t prof_used = 0;

    template <typename T> int dmalloc(T** p, size_t n) {
        CK(cudaMalloc((void**)p, n * sizeof(T)));
        (weight_scope ? w_allocs : allocs).push_back(*p);
        return 0;
    }
    template <typename T> int upload(T** p, const std::vector<T>& v) {
        if (dmalloc(p, v.size())) return 1;
        CK(cudaMemcpy(*p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
        return 0;
    }
};

// ---------------------------------------------------------------------------------------------
// parameter inventory (mirrors the module constructors; names are the reference state-dict keys)
// ---------------------------------------------------------------------------------------------
namespace {

struct Inventory {
    std::vector<std::pair<std::string, std::vector<int64_t>>> items;
    void conv(const std::string& n, int co, int ci, int k) {
        items.push_back({n + ".weight", {co, ci, k, k, k}});
        items.push_back({n + ".bias", {co}});
    }
    void lin(const std::string& n, int co, int ci) {
        items.push_back({n + ".weight", {co, ci}});
        items.push_back({n + ".bias", {co}});
    }
    void resblock(const std::string& p, int ci, int co, int E) {  // ResBlockModulateBEST :220-235
        conv(p + ".conv20", co, ci, 3);
        conv(p + ".conv21", co, co, 3);
        lin(p + ".dense1.dense", ci, E);
        lin(p + ".dense2.dense", co, E);
        if (ci != co) conv(p + ".res_conv", co, ci, 1);
    }
    void wavelet(const std::string& p, int c, int E, int hi) {  // WaveletUPorDown :334-367
        conv(p + ".Conv_0", c, c, 3);
        lin(p + ".Dense_0", c, E);
        conv(p + ".Conv_1", c, c, 3);
        conv(p + ".Conv_2", c, c, 1);
        if (hi > 0) {
            items.push_back({p + ".convH_0.0.weight", {3 * c, hi, 3, 3, 3}});
            items.push_back({p + ".convH_0.0.bias", {3 * c}});
        }
        lin(p + ".dense1.dense", c, E);
        lin(p + ".dense2.dense", c, E);
    }
};

Inventory build_inventory(const tmd_config& cfg) {
    const int* c = cfg.channels;
    const int E = cfg.embed_dim;
    Inventory I;
    I.lin("embed.0", E, cfg.inter_dim);
    I.lin("embed.2", E, E);
    I.lin("embed2.0", 4 * E, 768);
    I.lin("embed2.2", 4 * E, 4 * E);
    I.lin("embed2.4", E, 4 * E);
    for (const char* s : {"conv1", "conv2"}) {
        I.conv(std::string(s) + ".conv20", c[0], 1, 1);
        I.conv(std::string(s) + ".conv21", c[0], c[0], 3);
        I.lin(std::string(s) + ".dense2.dense", c[0], E);
    }
    const char* downs[6] = {"down1", "down2", "down3", "down1_1", "down2_1", "down3_1"};
    for (int i = 0; i < 6; ++i) {
        const int l = i % 3;
        I.resblock(std::string(downs[i]) + ".conv20", c[l], c[l + 1], E);
        I.wavelet(std::string(downs[i]) + ".down", c[l + 1], E, 0);
    }
    I.resblock("middle1", c[3], c[3], E);
    const char* ups[3] = {"up1", "up2", "up3"};
    for (int i = 0; i < 3; ++i) {
        const int ci = c[3 - i], co = c[2 - i];
        I.wavelet(std::string(ups[i]) + ".up1", co, E, ci);
        I.resblock(std::string(ups[i]) + ".conv20", 3 * ci, co, E);
    }
    for (int i = 0; i < 4; ++i) I.resblock("final.conv2" + std::to_string(i), i == 0 ? 3 * c[0] : c[0], c[0], E);
    I.conv("final.conv24", 1, c[0], 1);
    I.lin("final.dense2.dense", c[0], E);
    return I;
}

bool ends_with(const std::string& s, const std::string& suf) {
    return s.size() >= suf.size() && s.compare(s.size() - suf.size(), suf.size(), suf) == 0;
}
bool contains(const std::string& s, const std::string& sub) { return s.find(sub) != std::string::npos; }

}  // namespace

// ---------------------------------------------------------------------------------------------
// C ABI: lifecycle + weights
// ---------------------------------------------------------------------------------------------
extern "C" const char* tmd_last_error(void) { return g_err; }
int tmd::set_last_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return 1;
}
extern "C" int tmd_version(void) { return 1; }

extern "C" int tmd_create(const tmd_config* cfg, tmd_handle* out) {
    if (!cfg || !out) return fail("tmd_create: null argument");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return fail("tmd_create: no CUDA device (this library has no CPU path)");
    if (cfg->height % 8 || cfg->width % 8 || cfg->height <= 0 || cfg->width <= 0)
        return fail("tmd_create: H and W must be positive multiples of 8 (three Haar levels), got %dx%d", cfg->height, cfg->width);
    for (int i = 0; i < 4; ++i)
        if (cfg->channels[i] <= 0 || cfg->channels[i] % 8) return fail("tmd_create: channels[%d]=%d must be a positive multiple of 8", i, cfg->channels[i]);
    if (cfg->dtype != TMD_F32 && cfg->dtype != TMD_BF16) return fail("tmd_create: dtype must be TMD_F32 or TMD_BF16");
    if (cfg->max_batch <= 0 || cfg->max_bands <= 0) return fail("tmd_create: max_batch and max_bands must be positive");
    if (cfg->embed_dim % 32 || cfg->inter_dim % 2) return fail("tmd_create: embed_dim must be a multiple of 32 and inter_dim even");
    auto* e = new tmd_engine();
    e->cfg = *cfg;
    e->dtype = cfg->dtype;
    e->esize = cfg->dtype == TMD_F32 ? 4 : 2;
    cudaGetDevice(&e->device);
    *out = e;
    return 0;
}

extern "C" int tmd_destroy(tmd_handle h) {
    if (!h) return 0;
    for (auto* p : h->plans) conv_umma_plan_destroy(p);
    for (void* p : h->allocs) cudaFree(p);
    for (void* p : h->w_allocs) cudaFree(p);
    for (auto& sp : h->prof_events) { cudaEventDestroy(sp.a); cudaEventDestroy(sp.b); }
    if (h->arena) cudaFree(h->arena);
    h->drop_graphs();
    if (h->cap_stream) cudaStreamDestroy(h->cap_stream);
    delete h;
    return 0;
}

extern "C" int tmd_load_weight(tmd_handle h, const char* key, const float* host, const int64_t* shape, int32_t ndim) {
    if (!h || !key || !host || !shape) return fail("tmd_load_weight: null argument");
    auto& inv = h->inv_shapes;
    if (inv.empty())
        for (auto& it : build_inventory(h->cfg).items) inv[it.first] = it.second;
    auto it = inv.find(key);
    if (it == inv.end()) return fail("tmd_load_weight: unexpected key '%s'", key);
    std::vector<int64_t> shp(shape, shape + ndim);
    if (shp != it->second) return fail("tmd_load_weight: shape mismatch for '%s'", key);
    size_t n = 1;
    for (auto d : shp) n *= (size_t)d;
    HostTensor t;
    t.data.assign(host, host + n);
    t.shape = shp;
    h->host[key] = std::move(t);
    h->finalized = false;
    return 0;
}

extern "C" int tmd_set_time_frequencies(tmd_handle h, const float* host, int32_t n) {
    if (!h || !host) return fail("tmd_set_time_frequencies: null argument");
    if (n != h->cfg.inter_dim / 2) return fail("tmd_set_time_frequencies: expected %d frequencies", h->cfg.inter_dim / 2);
    h->freqs_host.assign(host, host + n);
    h->finalized = false;
    return 0;
}

static int pack_conv(tmd_engine* e, const std::string& name, const float* w, const float* bias, int Co, int Ci, int k, float bias_scale) {
    const int taps = k * k * k;
    ConvW cw;
    cw.Co = Co; cw.Ci = Ci; cw.taps = taps;
    std::vector<float> pf((size_t)taps * Ci * Co);
    std::vector<__nv_bfloat16> pb((size_t)taps * Ci * Co);
    for (int co = 0; co < Co; ++co)
        for (int ci = 0; ci < Ci; ++ci)
            for (int t = 0; t < taps; ++t) {
                const float v = w[((size_t)co * Ci + ci) * taps + t];  // [Co][Ci][kd][kh][kw], t = (kd*k+kh)*k+kw
                pf[((size_t)t * Ci + ci) * Co + co] = v;
                pb[(size_t)co * taps * Ci + (size_t)t * Ci + ci] = __float2bfloat16(v);
            }
    if (e->upload(&cw.w_f32, pf)) return 1;
    if (e->upload(&cw.w_bf16, pb)) return 1;
    if (bias) {
        std::vector<float> b(bias, bias + Co);
        for (auto& v : b) v *= bias_scale;
        if (e->upload(&cw.bias, b)) return 1;
    }
    e->convs[name] = cw;
    return 0;
}

extern "C" int tmd_finalize_weights(tmd_handle h) {
    if (!h) return fail("tmd_finalize_weights: null handle");
    auto inv = build_inventory(h->cfg);
    for (auto& it : inv.items)
        if (!h->host.count(it.first)) return fail("tmd_finalize_weights: missing weight '%s'", it.first.c_str());
    if (h->freqs_host.empty()) return fail("tmd_finalize_weights: tmd_set_time_frequencies was not called");
    CK(cudaSetDevice(h->device));
    // a re-finalize (second load_state_dict, checkpoint sweep) replaces every repacked tensor: drop what the previous one
    // uploaded, and the launch lists / tensor maps that point into it
    CK(cudaDeviceSynchronize());
    for (auto* p : h->plans) conv_umma_plan_destroy(p);
    h->plans.clear();
    h->pre_ops.clear();
    h->step_ops.clear();
    for (void* p : h->w_allocs) cudaFree(p);
    h->w_allocs.clear();
    struct Scope { tmd_engine* e; ~Scope() { e->weight_scope = false; } } scope{h};
    h->weight_scope = true;
    const int E = h->cfg.embed_dim;
    auto H = [&](const std::string& k) -> const std::vector<float>& { return h->host[k].data; };

    // ---- convolutions ----
    h->convs.clear();
    for (auto& it : inv.items) {
        const std::string& key = it.first;
        if (!ends_with(key, ".weight") || it.second.size() != 5) continue;
        const std::string name = key.substr(0, key.size() - 7);
        const int Co = (int)it.second[0], Ci = (int)it.second[1], k = (int)it.second[2];
        const float* w = H(key).data();
        const float* b = H(name + ".bias").data();
        // modulated convs never use their bias (callers pass bias=None, :171, :247, :271, :404)
        const bool modulated = ends_with(name, ".conv21") || ends_with(name, ".Conv_1") || name == "final.conv24";
        if (contains(name, ".convH_0.0")) {
            // groups=3 (:359-361): three independent convs; skipH = convH(cat/2)*2 == W*S + 2b (:381)
            const int co = Co / 3;
            for (int g = 0; g < 3; ++g)
                if (pack_conv(h, name + "#" + std::to_string(g), w + (size_t)g * co * Ci * 27, b + g * co, co, Ci, k, 2.0f)) return 1;
        } else if (Ci == 1 && k == 1) {
            const int s = name.rfind("conv1", 0) == 0 ? 0 : 1;   // stems conv1.conv20 / conv2.conv20
            std::vector<float> wv(w, w + Co), bv(b, b + Co);
            if (h->upload(&h->d_stem_w[s], wv) || h->upload(&h->d_stem_b[s], bv)) return 1;
        } else if (Co == 1 && k == 1) {
            std::vector<float> wv(w, w + Ci);
            if (h->upload(&h->d_proj_w, wv)) return 1;
        } else {
            if (pack_conv(h, name, w, modulated ? nullptr : b, Co, Ci, k, 1.0f)) return 1;
            // res_conv over a concat [h | condition skip | main skip] (up*.conv20, final.conv20): the middle third of its input is
            // step-invariant -> "#cond" (with the bias) runs once per batch, "#var" (the outer thirds) rides in conv21's K loop
            if (k == 1 && ends_with(name, ".res_conv") && Ci % 3 == 0 && (name.rfind("up", 0) == 0 || name.rfind("final", 0) == 0)) {
                const int c3 = Ci / 3;
                std::vector<float> wc((size_t)Co * c3), wv((size_t)Co * 2 * c3);
                for (int co = 0; co < Co; ++co)
                    for (int ci = 0; ci < Ci; ++ci) {
                        const float v = w[(size_t)co * Ci + ci];
                        if (ci >= c3 && ci < 2 * c3) wc[(size_t)co * c3 + ci - c3] = v;
                        else wv[(size_t)co * 2 * c3 + (ci < c3 ? ci : ci - c3)] = v;
                    }
                if (pack_conv(h, name + "#cond", wc.data(), b, Co, c3, 1, 1.0f)) return 1;
                if (pack_conv(h, name + "#var", wv.data(), nullptr, Co, 2 * c3, 1, 1.0f)) return 1;
            }
        }
    }

    // ---- conditioning tables ----
    h->time_off.clear(); h->text_off.clear();
    std::vector<float> Wt, bt, Ws, bs;
    auto append = [&](std::vector<float>& Wm, std::vector<float>& bm, std::map<std::string, int>& off, const std::string& mod,
                      const std::string& key) {
        const auto& w = H(key + ".weight");
        const auto& b = H(key + ".bias");
        off[mod] = (int)bm.size();
        Wm.insert(Wm.end(), w.begin(), w.end());
        bm.insert(bm.end(), b.begin(), b.end());
    };
    for (auto& it : inv.items) {
        const std::string& key = it.first;
        if (!ends_with(key, ".weight") || it.second.size() != 2) continue;
        const std::string name = key.substr(0, key.size() - 7);
        if (name.rfind("embed", 0) == 0) continue;
        const bool cond_branch = contains(name, "_1.");   // flag=True blocks never read their time Dense (:238-241, :400-401)
        const bool wavelet = contains(name, ".down.") || contains(name, ".up1.");
        if (ends_with(name, ".Dense_0")) {
            if (!cond_branch) append(Wt, bt, h->time_off, name, name);
        } else if (ends_with(name, ".dense1.dense")) {
            const std::string mod = name.substr(0, name.size() - 6);
            if (wavelet) append(Ws, bs, h->text_off, mod, name);             // s = dense1(zemb) (:404)
            else if (!cond_branch) append(Wt, bt, h->time_off, mod, name);   // h = x + dense1(embed) (:239)
        } else if (ends_with(name, ".dense2.dense")) {
            const std::string mod = name.substr(0, name.size() - 6);
            if (!wavelet) append(Ws, bs, h->text_off, mod, name);            // s = dense2(prompt) (:171, :247, :271); wavelet dense2 is unused
        }
    }
    h->time_rows = (int)bt.size();
    h->text_rows = (int)bs.size();
    if (h->upload(&h->d_Wt, Wt) || h->upload(&h->d_bt, bt) || h->upload(&h->d_Ws, Ws) || h->upload(&h->d_bs, bs)) return 1;
    if (h->upload(&h->d_freqs, h->freqs_host)) return 1;
    if (h->upload(&h->d_e_w1, H("embed.0.weight")) || h->upload(&h->d_e_b1, H("embed.0.bias")) ||
        h->upload(&h->d_e_w2, H("embed.2.weight")) || h->upload(&h->d_e_b2, H("embed.2.bias")))
        return 1;
    const char* pk[3] = {"embed2.0", "embed2.2", "embed2.4"};
    for (int i = 0; i < 3; ++i)
        if (h->upload(&h->d_p_w[i], H(std::string(pk[i]) + ".weight")) || h->upload(&h->d_p_b[i], H(std::string(pk[i]) + ".bias"))) return 1;
    const int MB = h->cfg.max_batch;
    if (!h->d_text) {
        h->weight_scope = false;      // sized by the configuration, not by the weights: kept across re-finalizes
        if (h->dmalloc(&h->d_text, (size_t)TMD_NUM_SENSORS * 768) || h->dmalloc(&h->d_p_tmp0, (size_t)TMD_NUM_SENSORS * 4 * E) ||
            h->dmalloc(&h->d_p_tmp1, (size_t)TMD_NUM_SENSORS * 4 * E) || h->dmalloc(&h->d_prompt, (size_t)TMD_NUM_SENSORS * E) ||
            h->dmalloc(&h->d_tin, (size_t)MB) || h->dmalloc(&h->d_embed, (size_t)MB * E) || h->dmalloc(&h->d_sensor, (size_t)MB))
            return 1;
        CK(cudaMemset(h->d_text, 0, (size_t)TMD_NUM_SENSORS * 768 * sizeof(float)));
        h->weight_scope = true;
    }
    if (h->dmalloc(&h->d_sc_sensor, (size_t)TMD_NUM_SENSORS * h->text_rows) || h->dmalloc(&h->d_tb, (size_t)MB * h->time_rows) ||
        h->dmalloc(&h->d_sc, (size_t)MB * h->text_rows) || h->dmalloc(&h->d_tb_all, (size_t)kMaxPlanEvals * h->time_rows) ||
        h->dmalloc(&h->d_tin_all, (size_t)kMaxPlanEvals))
        return 1;
    h->drop_graphs();
    h->text_dirty = true;
    h->finalized = true;
    h->B = 0;  // force a re-plan
    CK(cudaDeviceSynchronize());
    return 0;
}

extern "C" int tmd_set_text_embedding(tmd_handle h, int32_t sensor, const float* host768) {
    if (!h || !host768) return fail("tmd_set_text_embedding: null argument");
    if (sensor < 0 || sensor >= TMD_NUM_SENSORS) return fail("tmd_set_text_embedding: sensor id %d out of range", sensor);
    if (!h->finalized) return fail("tmd_set_text_embedding: call tmd_finalize_weights first");
    CK(cudaMemcpy(h->d_text + (size_t)sensor * 768, host768, 768 * sizeof(float), cudaMemcpyHostToDevice));
    h->have_text[sensor] = true;
    h->text_dirty = true;
    return 0;
}

// ---------------------------------------------------------------------------------------------
// plan: bind a geometry, carve the workspace, build the launch lists
// ---------------------------------------------------------------------------------------------
namespace {

struct Planner {
    tmd_engine* e;
    std::vector<Launch>* ops;
    int B, N, H, W;
    const char* err = nullptr;

    size_t vox(int lvl) const { return (size_t)B * N * (H >> lvl) * (W >> lvl); }
    void push(int kind, double flops, double bytes, LaunchFn fn, const std::string& label = "") {
        Launch l;
        l.fn = std::move(fn); l.kind = kind; l.flops = flops; l.bytes = bytes; l.label = label;
        ops->push_back(std::move(l));
    }

    View buf(const std::string& name, int lvl, int C) {
        auto it = e->bufs.find(name);
        if (it == e->bufs.end()) {
            const size_t bytes = (vox(lvl) * C * e->esize + 255) & ~(size_t)255;
            const size_t rz = e->redzone_on ? kRedzoneBytes : 0;
            if (e->arena_used + bytes + rz > e->arena_size) { err = "workspace arena exhausted"; return View{}; }
            void* p = e->arena + e->arena_used;
            e->arena_used += bytes + rz;
            if (rz) e->redzones.push_back({name, (char*)p + bytes});
            it = e->bufs.emplace(name, p).first;
        }
        return View{it->second, C, C};
    }
    View slice(View v, int c0, int C) const { return View{(char*)v.p + (size_t)c0 * e->esize, C, v.ld}; }
    const float* tb(const std::string& mod, int c0 = 0) {
        auto it = e->time_off.find(mod);
        if (it == e->time_off.end()) { err = "missing time table entry"; return nullptr; }
        return e->d_tb + it->second + c0;
    }
    const float* sc(const std::string& mod) {
        auto it = e->text_off.find(mod);
        if (it == e->text_off.end()) { err = "missing text table entry"; return nullptr; }
        return e->d_sc + it->second;
    }

    // bf16 mode: the 1x1x1 convs that are summed with a 3x3x3 conv's result (ResBlock res_conv, wavelet-down Conv_2)
    // run as extra K chunks of that conv (ConvOp::src2) instead of as launches of their own
    bool fuse = false;

    void conv(const std::string& wname, View src, int lvl, int mode, const OutSpec& out, View add0 = View{}, View add1 = View{},
              View* hf = nullptr, const std::string& w2name = "", View src2 = View{}, int gap_at = 0, int gap = 0) {
        auto it = e->convs.find(wname);
        if (it == e->convs.end()) { err = "missing conv weights"; return; }
        const ConvW& cw = it->second;
        if (cw.Ci != src.C) { err = "conv input channel mismatch"; return; }
        const ConvW* cw2 = nullptr;
        if (!w2name.empty()) {
            auto it2 = e->convs.find(w2name);
            if (it2 == e->convs.end()) { err = "missing conv weights (fused 1x1x1)"; return; }
            cw2 = &it2->second;
            if (cw2->taps != 1 || cw2->Co != cw.Co || cw2->Ci + gap != src2.C || (cw.bias && cw2->bias)) { err = "fused 1x1x1 conv mismatch"; return; }
        }
        ConvOp op{};
        op.B = B; op.N = N; op.H = H >> lvl; op.W = W >> lvl;
        op.Ci = cw.Ci; op.ld_in = src.ld; op.src = src.p;
        op.Co = cw.Co; op.taps = cw.taps;
        op.w_f32 = cw.w_f32; op.w_bf16 = cw.w_bf16;
        ConvEpilogue& ep = op.epi;
        ep.mode = mode;
        ep.bias = cw.bias;
        if (cw2) {
            op.src2 = src2.p; op.Ci2 = cw2->Ci; op.ld_in2 = src2.ld; op.w2_bf16 = cw2->w_bf16;
            op.src2_gap_at = gap_at; op.src2_gap = gap;
            if (cw2->bias) ep.bias = cw2->bias;
        }
        ep.pool_out = out.pool.p; ep.ld_pool = out.pool.ld;
        ep.add0 = add0.p; ep.ld_add0 = add0.ld;
        ep.raw_f16 = out.raw.f16 ? 1 : 0; ep.add0_f16 = add0.f16 ? 1 : 0;

        ep.add1 = add1.p; ep.ld_add1 = add1.ld;
        ep.raw_out = out.raw.p; ep.ld_raw = out.raw.ld;
        for (int i = 0; i < 2; ++i) {
            ep.act_out[i] = out.act[i].dst.p; ep.ld_act[i] = out.act[i].dst.ld;
            ep.tb[i] = out.act[i].tb; ep.sc[i] = out.act[i].sc;
        }
        ep.ld_tb = e->time_rows; ep.ld_sc = e->text_rows;
        if (out.proj) { ep.proj_out = out.proj; ep.proj_w = e->d_proj_w; }
        if (hf) { for (int i = 0; i < 3; ++i) ep.hf_out[i] = hf[i].p; ep.ld_hf = hf[0].ld; }
        // algorithmic cost of this launch: 2*MACs; every operand/result tensor touched once
        const double nv = (double)vox(lvl);
        const double flops = 2.0 * nv * cw.Ci * cw.taps * cw.Co + (cw2 ? 2.0 * nv * cw2->Ci * cw.Co : 0.0);
        const double ovox = mode == CONV_POOL ? nv / 4 : (mode == CONV_UPSAMPLE ? nv * 4 : nv);
        int n_out = (out.raw.p ? 1 : 0) + (out.act[0].dst.p ? 1 : 0) + (out.act[1].dst.p ? 1 : 0);
        double bytes = nv * cw.Ci * e->esize + ovox * cw.Co * e->esize * n_out + (double)cw.Ci * cw.taps * cw.Co * e->esize;
        if (add0.p) bytes += ovox * cw.Co * e->esize;
        if (add1.p) bytes += ovox / 4 * cw.Co * e->esize;
        if (hf) bytes += 3 * ovox * cw.Co * e->esize;
        if (cw2) bytes += nv * cw2->Ci * e->esize + (double)cw2->Ci * cw.Co * e->esize;
        if (out.pool.p) bytes += nv / 4 * cw.Co * e->esize;
        const int kind = cw.taps == 27 ? TMD_K_CONV3 : TMD_K_CONV1;
        // issued FLOPs: the tcgen05 kernel neither loads nor multiplies the band planes outside [0, N) (zero padding of the
        // band axis): of the 3N (plane, dz) pairs of a 3x3x3 conv 3N - 2 remain.  The algorithmic count above includes them,
        // as the reference executes them.
        double flops_exec = flops;
        if (e->dtype == TMD_BF16 && cw.taps == 27 && conv_umma_supported(op) && conv_umma_skips_padding_planes(op))
            flops_exec = 2.0 * nv * cw.Ci * cw.taps * cw.Co * (3.0 * N - 2.0) / (3.0 * N) + (cw2 ? 2.0 * nv * cw2->Ci * cw.Co : 0.0);
        char lb[160];
        snprintf(lb, sizeof(lb), "%s Ci=%d Co=%d L%d %s", wname.c_str(), cw.Ci, cw.Co, lvl,
                 mode == CONV_POOL ? "pool" : (mode == CONV_UPSAMPLE ? "up" : "same"));
        const std::string label = std::string(lb) + (cw2 ? " +" + w2name.substr(w2name.rfind('.') + 1) : "") + (out.pool.p ? " +pool" : "");
        if ((cw2 || out.pool.p || ep.raw_f16 || ep.add0_f16 || ep.proj_out) && (e->dtype == TMD_F32 || !conv_umma_supported(op))) { err = "fused conv outside the tcgen05 envelope"; return; }
        if (e->dtype == TMD_F32) {
            push(kind, flops, bytes, [op](cudaStream_t s) { return conv_simt_f32(op, s); }, label);
        } else if (conv_umma_supported(op)) {
            UmmaPlan* plan = conv_umma_plan_create(op);
            if (!plan) { err = "tensor map creation failed"; return; }
            e->plans.push_back(plan);
            push(kind, flops, bytes, [op, plan](cudaStream_t s) { return conv_umma_bf16(op, plan, s); }, label + " umma");
            ops->back().flops_exec = flops_exec;
        } else {
            push(kind, flops, bytes, [op](cudaStream_t s) { return conv_simt_bf16(op, s); }, label + " simt");
        }
    }

    // ResBlockModulateBEST.forward (:237-249).  in_act = swish(x [+ dense1(embed)]) prepared by the producer.
    // rc (fused plans, concat inputs [h | condition skip | main skip]): res_conv of the condition slice + bias, computed once per
    // batch by the step-invariant pass; the K chunks of the fused res_conv then skip that slice and rc arrives through add0.
    void resblock(const std::string& p, View in_act, View in_raw, int Co, int lvl, const OutSpec& out, View rc = View{}) {
        View a2 = buf("a2@" + std::to_string(lvl), lvl, cmax(lvl));
        a2.C = Co;
        OutSpec o1;
        o1.act[0] = ActSpec{a2, nullptr, sc(p + ".dense2")};
        conv(p + ".conv20", in_act, lvl, CONV_NORMAL, o1);
        View res = in_raw;
        if (fuse && rc.p && e->convs.count(p + ".res_conv#var")) {
            const int third = in_raw.C / 3;
            conv(p + ".conv21", a2, lvl, CONV_NORMAL, out, rc, View{}, nullptr, p + ".res_conv#var", in_raw, third, third);
            return;
        }
        if (fuse && e->convs.count(p + ".res_conv")) {
            conv(p + ".conv21", a2, lvl, CONV_NORMAL, out, View{}, View{}, nullptr, p + ".res_conv", in_raw);
            return;
        }
        if (e->convs.count(p + ".res_conv")) {
            View r = buf("r@" + std::to_string(lvl), lvl, cmax(lvl));
            r.C = Co;
            OutSpec o2;
            o2.raw = r;
            conv(p + ".res_conv", in_raw, lvl, CONV_NORMAL, o2);
            res = r;
        }
        conv(p + ".conv21", a2, lvl, CONV_NORMAL, out, res);
    }

    // WaveletUPorDown.forward, down=True (:369-414); lvl = input level, outputs at lvl+1
    // in_pool (fused plans): 2x2 average of in_raw, emitted by its producer; Conv_2(pool(x)) == pool(Conv_2(x)).
    void wavelet_down(const std::string& p, View in_act, View in_raw, int C, int lvl, bool flag, View* hf, const OutSpec& out,
                      View in_pool = View{}) {
        View xp;
        if (!in_pool.p) {
            xp = buf("xp@" + std::to_string(lvl + 1), lvl + 1, cmax(lvl + 1));
            xp.C = C;
            OutSpec o1;
            o1.raw = xp;
            conv(p + ".Conv_2", in_raw, lvl, CONV_POOL, o1);
        }
        View ac1 = buf("ac1@" + std::to_string(lvl + 1), lvl + 1, cmax(lvl + 1));
        ac1.C = C;
        OutSpec o2;
        o2.act[0] = ActSpec{ac1, flag ? nullptr : tb(p + ".Dense_0"), sc(p + ".dense1")};
        conv(p + ".Conv_0", in_act, lvl, CONV_POOL, o2, View{}, View{}, hf);
        if (in_pool.p) conv(p + ".Conv_1", ac1, lvl + 1, CONV_NORMAL, out, View{}, View{}, nullptr, p + ".Conv_2", in_pool);
        else conv(p + ".Conv_1", ac1, lvl + 1, CONV_NORMAL, out, xp);
    }

    // WaveletUPorDown.forward, up=True; lvl = input (low) level, outputs at lvl-1.  T = IDWT(0, skipH').
    void wavelet_up(const std::string& p, View in_act, View in_raw, int C, int lvl, View T, const OutSpec& out) {
        View xl = buf("xl@" + std::to_string(lvl), lvl, cmax(lvl));
        xl.C = C;
        OutSpec o1;
        o1.raw = xl;
        conv(p + ".Conv_2", in_raw, lvl, CONV_NORMAL, o1);
        View ac1 = buf("ac1u@" + std::to_string(lvl - 1), lvl - 1, cmax(lvl - 1));
        ac1.C = C;
        OutSpec o2;
        o2.act[0] = ActSpec{ac1, tb(p + ".Dense_0"), sc(p + ".dense1")};
        conv(p + ".Conv_0", in_act, lvl, CONV_UPSAMPLE, o2, T);
        conv(p + ".Conv_1", ac1, lvl - 1, CONV_NORMAL, out, T, xl);
    }

    int cmax(int lvl) const { return e->cfg.channels[lvl < 3 ? lvl + 1 : 3]; }
};

int build_plan(tmd_engine* e, int B, int N, int H, int W) {
    const int* c = e->cfg.channels;
    for (auto* p : e->plans) conv_umma_plan_destroy(p);
    e->plans.clear();
    e->drop_graphs();
    e->pre_ops.clear();
    e->step_ops.clear();
    e->bufs.clear();
    e->redzones.clear();
    e->arena_used = 0;
    {
        const char* rz = getenv("TMD_ARENA_REDZONE");
        e->redzone_on = rz && rz[0] == '1';
    }
    if (!e->arena) {
        // worst-case footprint at (max_batch, max_bands, H, W): sum over levels of channel-widths of all named buffers
        const size_t v0 = (size_t)e->cfg.max_batch * e->cfg.max_bands * e->cfg.height * e->cfg.width;
        size_t elems = 0;
        const int cm[4] = {c[1], c[2], c[3], c[3]};
        for (int l = 0; l < 4; ++l) {
            const size_t v = v0 >> (2 * l);
            elems += v * (size_t)(2 * 3 * c[l] + 8 * cm[l] + 8 * c[l]) + v / 4 * (size_t)cm[l];
        }
        e->arena_size = elems * e->esize + ((size_t)64 << 20) + (e->redzone_on ? 256 * kRedzoneBytes : 0);
        CK(cudaMalloc((void**)&e->arena, e->arena_size));
        for (int i = 0; i < TMD_NUM_SLOTS; ++i)
            if (e->dmalloc(&e->slots[i], v0)) return 1;
        if (e->dmalloc(&e->d_out, v0) || e->dmalloc(&e->d_xdup, v0)) return 1;
        char* ws = nullptr;
        if (e->dmalloc(&ws, threshold_workspace_bytes(e->cfg.max_batch))) return 1;
        e->d_th_ws = ws;
        CK(cudaMemset(ws, 0, threshold_workspace_bytes(e->cfg.max_batch)));   // each threshold call leaves it clean again
    }
    Planner P{e, &e->pre_ops, B, N, H, W};
    {
        bool ok = e->dtype == TMD_BF16 && !getenv("TMD_NO_FUSE_1X1");
        for (int l = 0; l < 4; ++l) ok = ok && c[l] % 32 == 0;
        ConvOp probe{};
        probe.B = B; probe.N = N; probe.H = H >> 2; probe.W = W >> 2; probe.Ci = probe.ld_in = 32; probe.Co = 32; probe.taps = 27;
        P.fuse = ok && conv_umma_supported(probe);
    }
    const int dt = e->dtype;
    const int tr = e->time_rows, sr = e->text_rows;
    const int HW = H * W;
    auto L = [](int l) { return std::to_string(l); };

    // concat buffers: [h | cond skip | main skip] per level (:631-634), raw and activated forms
    View cat_raw[4], cat_act[4];
    for (int l = 0; l < 4; ++l) {
        cat_raw[l] = P.buf("cat_raw@" + L(l), l, 3 * c[l]);
        cat_act[l] = P.buf("cat_act@" + L(l), l, 3 * c[l]);
    }
    View Tup[3];  // IDWT(0, skipH') for up1 (level 2), up2 (level 1), up3 (level 0)
    for (int l = 0; l < 3; ++l) Tup[l] = P.buf("T@" + L(l), l, c[l]);

    // ======================= step-invariant condition branch (:611-618) =======================
    {
        View a_stem = P.buf("a_stem", 0, c[0]);
        const float* scs = P.sc("conv1.dense2");
        P.push(TMD_K_ELEMENTWISE, 0, (double)P.vox(0) * (4 + c[0] * e->esize), [=](cudaStream_t s) {
            return stem_expand(dt, e->cur_ms, e->cur_pan, e->d_stem_w[0], e->d_stem_b[0], scs, sr, a_stem.p, c[0], B, N, HW, s);
        }, "stem_expand (condition)");
        View in_act = P.buf("in_act@0", 0, c[0]);
        OutSpec o;
        o.raw = P.slice(cat_raw[0], c[0], c[0]);
        o.act[0] = ActSpec{in_act, nullptr, nullptr};
        P.conv("conv1.conv21", a_stem, 0, CONV_NORMAL, o);
        View in_raw = o.raw;
        const char* names[3] = {"down1_1", "down2_1", "down3_1"};
        for (int l = 0; l < 3; ++l) {
            const std::string nm = names[l];
            View y = P.buf("y@" + L(l), l, c[l + 1]);
            View ay = P.buf("ay@" + L(l), l, c[l + 1]);
            OutSpec ro;
            ro.raw = y;
            ro.act[0] = ActSpec{ay, nullptr, nullptr};
            View yq;
            if (P.fuse) ro.pool = yq = P.buf("yq@" + L(l + 1), l + 1, c[l + 1]);
            P.resblock(nm + ".conv20", in_act, in_raw, c[l + 1], l, ro);
            View hf[3];
            for (int g = 0; g < 3; ++g) hf[g] = P.buf("S" + L(g) + "@" + L(l + 1), l + 1, c[l + 1]);
            View nxt = P.buf("in_act@" + L(l + 1), l + 1, c[l + 1]);
            OutSpec wo;
            wo.raw = P.slice(cat_raw[l + 1], c[l + 1], c[l + 1]);
            if (l < 2) wo.act[0] = ActSpec{nxt, nullptr, nullptr};
            P.wavelet_down(nm + ".down", ay, y, c[l + 1], l, true, hf, wo, yq);
            in_act = nxt;
            in_raw = wo.raw;
            // convH_0 of the matching up block consumes only these skips (:381): hoist it here
            const std::string up = std::string("up") + L(3 - l) + ".up1.convH_0.0#";
            View hq[3];
            for (int g = 0; g < 3; ++g) {
                hq[g] = P.buf("HQ" + L(g) + "@" + L(l + 1), l + 1, c[l]);
                OutSpec ho;
                ho.raw = hq[g];
                P.conv(up + L(g), hf[g], l + 1, CONV_NORMAL, ho);
            }
            View T = Tup[l];
            const size_t planes = (size_t)B * N;
            const int hh = H >> (l + 1), ww = W >> (l + 1), C = c[l];
            // IDWT: 3 coefficient planes in, one double-resolution tensor out
            P.push(TMD_K_ELEMENTWISE, 0, (double)planes * hh * ww * C * e->esize * 7.0,
                   [=](cudaStream_t s) { return idwt_highbands(dt, hq[0].p, hq[1].p, hq[2].p, T.p, C, planes, hh, ww, s); }, "idwt_highbands");
        }
    }

    // res_conv of the condition slices of the four concat inputs (up1/up2/up3.conv20, final.conv20): step-invariant too
    View Rc[4];
    {
        static const bool rc_env = !getenv("TMD_NO_RC_SPLIT");
        const char* cons[4] = {"final.conv20", "up3.conv20", "up2.conv20", "up1.conv20"};
        for (int l = 0; l < 4 && P.fuse && rc_env; ++l) {
            const std::string w = std::string(cons[l]) + ".res_conv#cond";
            if (!e->convs.count(w)) continue;
            const int Co = e->convs[w].Co;
            Rc[l] = P.buf("Rc@" + L(l), l, Co);
            Rc[l].f16 = true;                       // add-only tensor
            OutSpec ro;
            ro.raw = Rc[l];
            P.conv(w, P.slice(cat_raw[l], c[l], c[l]), l, CONV_NORMAL, ro);
        }
    }

    // ================================ per-step main branch ===================================
    P.ops = &e->step_ops;
    {
        const int E = e->cfg.embed_dim;
        P.push(TMD_K_COND, 0, ((double)tr * E + (double)B * tr) * 4, [=](cudaStream_t s) {
            if (e->tb_bcast_src) return broadcast_rows(e->tb_bcast_src, e->d_tb, B, tr, s);
            return time_table(e->cur_t_dev, e->cur_t_scalar, e->d_freqs, e->cfg.inter_dim, e->d_e_w1, e->d_e_b1, e->d_e_w2, e->d_e_b2,
                              e->d_Wt, e->d_bt, e->d_embed, e->d_tb, B, E, tr, s);
        }, "time_table");
        // activations of the cached condition slices for this step's time embedding
        const char* consumer[4] = {"final.conv20", "up3.conv20", "up2.conv20", "up1.conv20"};
        static const bool multi_env = !getenv("TMD_NO_SLICE_MULTI");
        bool multi = e->dtype == TMD_BF16 && multi_env;
        for (int l = 0; l < 4; ++l) multi = multi && c[l] % 8 == 0;
        if (multi) {
            struct Args { const void* src[4]; void* dst[4]; const float* tb[4]; int ld_src[4], ld_dst[4], C[4]; size_t vps[4]; } A{};
            double bytes = 0;
            for (int l = 0; l < 4; ++l) {
                View src = P.slice(cat_raw[l], c[l], c[l]), dst = P.slice(cat_act[l], c[l], c[l]);
                A.src[l] = src.p; A.dst[l] = dst.p; A.ld_src[l] = src.ld; A.ld_dst[l] = dst.ld; A.C[l] = c[l];
                A.tb[l] = P.tb(std::string(consumer[l]) + ".dense1", c[l]);
                A.vps[l] = (size_t)N * (H >> l) * (W >> l);
                bytes += 2.0 * B * A.vps[l] * c[l] * e->esize;
            }
            P.push(TMD_K_ELEMENTWISE, 0, bytes,
                   [=](cudaStream_t s) { return slice_activate_multi(4, A.src, A.ld_src, A.dst, A.ld_dst, A.tb, tr, A.C, B, A.vps, s); }, "slice_activate L0-L3");
        }
        for (int l = 0; l < 4 && !multi; ++l) {
            View src = P.slice(cat_raw[l], c[l], c[l]), dst = P.slice(cat_act[l], c[l], c[l]);
            const float* t = P.tb(std::string(consumer[l]) + ".dense1", c[l]);
            const int C = c[l];
            const size_t vps = (size_t)N * (H >> l) * (W >> l);
            P.push(TMD_K_ELEMENTWISE, 0, 2.0 * B * vps * C * e->esize,
                   [=](cudaStream_t s) { return slice_activate(dt, src.p, src.ld, dst.p, dst.ld, t, tr, C, B, vps, s); }, "slice_activate L" + L(l));
        }
        View a_stem = P.buf("a_stem", 0, c[0]);
        const float* scs = P.sc("conv2.dense2");
        P.push(TMD_K_ELEMENTWISE, 0, (double)P.vox(0) * (4 + c[0] * e->esize), [=](cudaStream_t s) {
            return stem_expand(dt, e->cur_x, nullptr, e->d_stem_w[1], e->d_stem_b[1], scs, sr, a_stem.p, c[0], B, N, HW, s);
        }, "stem_expand");
        View in_act = P.buf("in_act@0", 0, c[0]);
        OutSpec o;
        o.raw = P.slice(cat_raw[0], 2 * c[0], c[0]);
        o.act[0] = ActSpec{in_act, P.tb("down1.conv20.dense1"), nullptr};
        o.act[1] = ActSpec{P.slice(cat_act[0], 2 * c[0], c[0]), P.tb("final.conv20.dense1", 2 * c[0]), nullptr};
        P.conv("conv2.conv21", a_stem, 0, CONV_NORMAL, o);
        View in_raw = o.raw;
        const char* names[3] = {"down1", "down2", "down3"};
        const char* next_res[3] = {"down2.conv20", "down3.conv20", "middle1"};
        for (int l = 0; l < 3; ++l) {
            const std::string nm = names[l];
            View y = P.buf("y@" + L(l), l, c[l + 1]);
            View ay = P.buf("ay@" + L(l), l, c[l + 1]);
            OutSpec ro;
            ro.raw = y;
            ro.act[0] = ActSpec{ay, nullptr, nullptr};
            View yq;
            if (P.fuse) ro.pool = yq = P.buf("yq@" + L(l + 1), l + 1, c[l + 1]);
            P.resblock(nm + ".conv20", in_act, in_raw, c[l + 1], l, ro);
            View nxt = P.buf("in_act@" + L(l + 1), l + 1, c[l + 1]);
            OutSpec wo;
            wo.raw = P.slice(cat_raw[l + 1], 2 * c[l + 1], c[l + 1]);
            wo.act[0] = ActSpec{nxt, P.tb(std::string(next_res[l]) + ".dense1"), nullptr};
            wo.act[1] = ActSpec{P.slice(cat_act[l + 1], 2 * c[l + 1], c[l + 1]),
                                P.tb(std::string(consumer[l + 1]) + ".dense1", 2 * c[l + 1]), nullptr};
            P.wavelet_down(nm + ".down", ay, y, c[l + 1], l, false, nullptr, wo, yq);  // high bands of the main branch are never read (:620-627)
            in_act = nxt;
            in_raw = wo.raw;
        }
        {   // middle1 (:629)
            OutSpec mo;
            mo.raw = P.slice(cat_raw[3], 0, c[3]);
            mo.act[0] = ActSpec{P.slice(cat_act[3], 0, c[3]), P.tb("up1.conv20.dense1", 0), nullptr};
            P.resblock("middle1", in_act, in_raw, c[3], 3, mo);
        }
        for (int i = 0; i < 3; ++i) {  // up1..up3 (:631-633): level 3-i -> 2-i
            const int l = 3 - i;
            const std::string nm = std::string("up") + L(i + 1);
            View y = P.buf("y@" + L(l), l, P.cmax(l));
            y.C = c[l - 1];
            View ay = P.buf("ay@" + L(l), l, P.cmax(l));
            ay.C = c[l - 1];
            OutSpec ro;
            ro.raw = y;
            ro.act[0] = ActSpec{ay, nullptr, nullptr};
            P.resblock(nm + ".conv20", cat_act[l], cat_raw[l], c[l - 1], l, ro, Rc[l]);
            OutSpec wo;
            wo.raw = P.slice(cat_raw[l - 1], 0, c[l - 1]);
            wo.act[0] = ActSpec{P.slice(cat_act[l - 1], 0, c[l - 1]), P.tb(std::string(consumer[l - 1]) + ".dense1", 0), nullptr};
            P.wavelet_up(nm + ".up1", ay, y, c[l - 1], l, Tup[l - 1], wo);
        }
        {   // final (:265-273)
            View f[2] = {P.buf("f0", 0, c[0]), P.buf("f1", 0, c[0])};
            // the residual stream of the final block is only ever added (identity skips, :249): IEEE half instead of bf16
            static const bool f16_res = !getenv("TMD_NO_F16_RESIDUAL");
            if (P.fuse && f16_res) f[0].f16 = f[1].f16 = true;
            View af[2] = {P.buf("af0", 0, c[0]), P.buf("af1", 0, c[0])};
            View ia = cat_act[0], ir = cat_raw[0];
            static const bool proj_env = !getenv("TMD_NO_FUSE_PROJ");
            const bool proj_fused = P.fuse && proj_env;
            for (int i = 0; i < 4; ++i) {
                OutSpec fo;
                if (i < 3) {
                    fo.raw = f[i & 1];
                    fo.act[0] = ActSpec{af[i & 1], P.tb("final.conv2" + L(i + 1) + ".dense1"), nullptr};
                } else {
                    fo.act[0] = ActSpec{af[i & 1], nullptr, P.sc("final.dense2")};  // act(h) then conv24's modulation
                    if (proj_fused) {
                        // final.conv24 (1x1x1, 32 -> 1, :271-272) rides in this epilogue: the modulated activation never goes to HBM
                        fo.act[0].dst = View{};
                        fo.proj = e->d_out;
                        const size_t bytes = P.vox(0) * sizeof(float);
                        P.push(TMD_K_ELEMENTWISE, 0, (double)bytes, [=](cudaStream_t s) { return cudaMemsetAsync(e->d_out, 0, bytes, s); }, "zero projection target");
                        P.ops->back().is_kernel = false;
                    }
                }
                P.resblock("final.conv2" + L(i), ia, ir, c[0], 0, fo, i == 0 ? Rc[0] : View{});
                ia = af[i & 1];
                ir = f[i & 1];
            }
            const size_t nvox = P.vox(0);
            const int C = c[0];
            View last = af[1];
            if (proj_fused)      // the projection landed in the engine's own buffer: hand it to the caller's tensor where that is another one
                P.push(TMD_K_ELEMENTWISE, 0, 0.0, [=](cudaStream_t s) {
                    return e->cur_out == e->d_out ? cudaSuccess : cudaMemcpyAsync(e->cur_out, e->d_out, nvox * sizeof(float), cudaMemcpyDeviceToDevice, s);
                }, "network output");
            if (proj_fused) P.ops->back().is_kernel = false;
            else
                P.push(TMD_K_ELEMENTWISE, 2.0 * nvox * C, (double)nvox * (C * e->esize + 4),
                       [=](cudaStream_t s) { return project_to_band(dt, last.p, last.ld, e->d_proj_w, e->cur_out, C, nvox, s); }, "project_to_band");
        }
    }
    if (P.err) return fail("plan: %s", P.err);
    for (auto& rz : e->redzones) CK(cudaMemset(rz.second, 0xCB, kRedzoneBytes));
    e->B = B; e->N = N; e->H = H; e->W = W;
    return 0;
}

// Profiling (tmd_profile_begin/end): every launch is bracketed by a CUDA event pair on the launching
// stream; the pairs are resolved into per-kind totals at tmd_profile_end.
int prof_mark(tmd_engine* e, int kind, double flops, double bytes, cudaStream_t s, bool begin, const char* label = "", double flops_exec = -1.0) {
    if (!e->profiling) return 0;
    if (begin) {
        if (e->prof_used == e->prof_events.size()) {
            ProfSpan sp{};
            CK(cudaEventCreate(&sp.a));
            CK(cudaEventCreate(&sp.b));
            e->prof_events.push_back(sp);
        }
        ProfSpan& sp = e->prof_events[e->prof_used];
        sp.kind = kind; sp.flops = flops; sp.bytes = bytes; sp.label = label;
        sp.flops_exec = flops_exec < 0 ? flops : flops_exec;
        CK(cudaEventRecord(sp.a, s));
    } else {
        CK(cudaEventRecord(e->prof_events[e->prof_used].b, s));
        ++e->prof_used;
    }
    return 0;
}

int run_ops(tmd_engine* e, std::vector<Launch>& ops, cudaStream_t s) {
    for (auto& op : ops) {
        if (prof_mark(e, op.kind, op.flops, op.bytes, s, true, op.label.c_str(), op.flops_exec)) return 1;
        cudaError_t err = op(s);
        if (err != cudaSuccess) return fail("kernel launch failed: %s", cudaGetErrorString(err));
        if (prof_mark(e, op.kind, op.flops, op.bytes, s, false)) return 1;
        if (op.is_kernel) ++e->launches;
    }
    return 0;
}

int refresh_text(tmd_engine* e, cudaStream_t s) {
    const int E = e->cfg.embed_dim;
    // prompt = act(embed2(clip_emb)) (:602-603), then every text Dense
    CK(dense_rows(e->d_text, e->d_p_w[0], e->d_p_b[0], e->d_p_tmp0, TMD_NUM_SENSORS, 768, 4 * E, 1, s));
    CK(dense_rows(e->d_p_tmp0, e->d_p_w[1], e->d_p_b[1], e->d_p_tmp1, TMD_NUM_SENSORS, 4 * E, 4 * E, 1, s));
    CK(dense_rows(e->d_p_tmp1, e->d_p_w[2], e->d_p_b[2], e->d_prompt, TMD_NUM_SENSORS, 4 * E, E, 1, s));
    CK(dense_rows(e->d_prompt, e->d_Ws, e->d_bs, e->d_sc_sensor, TMD_NUM_SENSORS, E, e->text_rows, 0, s));
    e->launches += 4;
    e->text_dirty = false;
    return 0;
}

}  // namespace

extern "C" int tmd_precompute_condition(tmd_handle h, const float* pan, const float* ms, const int32_t* sensor_ids, int32_t B,
                                        int32_t bands, int32_t H, int32_t W, void* stream) {
    if (!h || !pan || !ms || !sensor_ids) return fail("tmd_precompute_condition: null argument");
    if (!h->finalized) return fail("tmd_precompute_condition: weights not finalized");
    if (B <= 0 || B > h->cfg.max_batch) return fail("tmd_precompute_condition: batch %d outside [1, %d]", B, h->cfg.max_batch);
    if (bands <= 0 || bands > h->cfg.max_bands) return fail("tmd_precompute_condition: bands %d outside [1, %d]", bands, h->cfg.max_bands);
    if (H % 8 || W % 8 || H <= 0 || W <= 0 || (size_t)H * W > (size_t)h->cfg.height * h->cfg.width)
        return fail("tmd_precompute_condition: patch %dx%d must be a multiple of 8 and fit the configured %dx%d", H, W, h->cfg.height, h->cfg.width);
    for (int i = 0; i < B; ++i) {
        if (sensor_ids[i] < 0 || sensor_ids[i] >= TMD_NUM_SENSORS) return fail("tmd_precompute_condition: sensor id %d out of range", sensor_ids[i]);
        if (!h->have_text[sensor_ids[i]]) return fail("tmd_precompute_condition: no text embedding set for sensor %d", sensor_ids[i]);
    }
    cudaStream_t s = (cudaStream_t)stream;
    CK(cudaSetDevice(h->device));
    if (B != h->B || bands != h->N || H != h->H || W != h->W)
        if (build_plan(h, B, bands, H, W)) return 1;
    if (h->text_dirty && refresh_text(h, s)) return 1;
    CK(cudaMemcpyAsync(h->d_sensor, sensor_ids, (size_t)B * sizeof(int), cudaMemcpyHostToDevice, s));
    CK(gather_rows(h->d_sc_sensor, h->d_sensor, h->d_sc, B, h->text_rows, s));
    ++h->launches;
    h->cur_pan = pan;
    h->cur_ms = ms;
    return run_ops(h, h->pre_ops, s);
}

static int denoise_impl(tmd_handle h, const float* x, const float* t_dev, float t_scalar, float* out, cudaStream_t s) {
    h->cur_x = x;
    h->cur_out = out;
    h->cur_t_dev = t_dev;
    h->cur_t_scalar = t_scalar;
    return run_ops(h, h->step_ops, s);
}

extern "C" int tmd_denoise(tmd_handle h, const float* x_t, const float* t_input_dev, float* out, void* stream) {
    if (!h || !x_t || !t_input_dev || !out) return fail("tmd_denoise: null argument");
    if (h->B == 0 || h->step_ops.empty()) return fail("tmd_denoise: call tmd_precompute_condition first");
    cudaStream_t s = (cudaStream_t)stream;
    CK(cudaSetDevice(h->device));
    return denoise_impl(h, x_t, t_input_dev, 0.f, out, s);
}

// The launch sequence of one plan: every evaluation's time table in one launch, then per step either one denoiser
// evaluation + x0 conversion / dynamic threshold, or one solver update.  Touches engine-owned buffers only (unless a
// trace is requested), so it can be captured into a CUDA graph.
static int sample_body(tmd_handle h, const tmd_plan* plan, float* trace, float* trace_x, int n_evals, bool pre_tb, cudaStream_t s) {
    static const bool fuse_updates = !getenv("TMD_NO_FUSE_UPDATE");
    const bool cfg = plan->cfg != 0;
    const bool noise_type = plan->model_type == TMD_MODEL_NOISE;
    const int Bx = cfg ? h->B / 2 : h->B;      // patches being sampled
    const size_t n_per = (size_t)h->N * h->H * h->W, n = n_per * Bx, n_net = n_per * h->B;
    auto slot_ok = [](int i) { return i >= 0 && i < TMD_NUM_SLOTS; };
    if (pre_tb) {
        if (prof_mark(h, TMD_K_COND, 0, ((double)h->time_rows * h->cfg.embed_dim + (double)n_evals * h->time_rows) * 4, s, true, "time_table_all"))
            return 1;
        CK(time_table(h->d_tin_all, 0.f, h->d_freqs, h->cfg.inter_dim, h->d_e_w1, h->d_e_b1, h->d_e_w2, h->d_e_b2, h->d_Wt, h->d_bt,
                      nullptr, h->d_tb_all, n_evals, h->cfg.embed_dim, h->time_rows, s));
        if (prof_mark(h, TMD_K_COND, 0, 0, s, false)) return 1;
        ++h->launches;
    }
    struct Unbind { tmd_engine* e; ~Unbind() { e->tb_bcast_src = nullptr; } } unbind{h};
    int n_eval = 0;
    for (int i = 0; i < plan->n_steps; ++i) {
        const tmd_step& st = plan->steps[i];
        if (st.kind == 0) {
            const tmd_eval& ev = st.eval;
            if (!slot_ok(ev.src) || !slot_ok(ev.dst)) return fail("tmd_sample: bad slot in step %d", i);
            float* net_out = trace ? trace + (size_t)n_eval * n_net : h->d_out;
            const float* net_in = h->slots[ev.src];
            if (trace_x) CK(cudaMemcpyAsync(trace_x + (size_t)n_eval * n, net_in, n * sizeof(float), cudaMemcpyDeviceToDevice, s));
            if (cfg) {   // x_in = cat([x] * 2)  (:338)
                CK(cudaMemcpyAsync(h->d_xdup, net_in, n * sizeof(float), cudaMemcpyDeviceToDevice, s));
                CK(cudaMemcpyAsync(h->d_xdup + n, net_in, n * sizeof(float), cudaMemcpyDeviceToDevice, s));
                net_in = h->d_xdup;
            }
            h->tb_bcast_src = pre_tb ? h->d_tb_all + (size_t)n_eval * h->time_rows : nullptr;
            if (denoise_impl(h, net_in, nullptr, ev.t_input, net_out, s)) return 1;
            // the singlestep update that consumes this x0 rides in the threshold's last pass (north_star item 5: the solver
            // update fused behind the evaluation; the per-sample quantile is why it cannot sit in the conv epilogue itself)
            FusedUpdate fu{};
            bool fused = false;
            if (fuse_updates && i + 1 < plan->n_steps && plan->steps[i + 1].kind == 1 && plan->steps[i + 1].update.form == 0) {
                const tmd_update& u = plan->steps[i + 1].update;
                if (slot_ok(u.dst) && slot_ok(u.x) && slot_ok(u.m) && u.m1 < TMD_NUM_SLOTS && u.dst != ev.dst) {
                    fu.dst = h->slots[u.dst]; fu.x = h->slots[u.x]; fu.m = h->slots[u.m]; fu.m1 = u.m1 >= 0 ? h->slots[u.m1] : nullptr;
                    fu.c1 = u.c1; fu.c2 = u.c2; fu.c3 = u.c3;
                    fused = true;
                }
            }
            // x, out in; x0 written, re-read three times (L2), re-written; fused update: its operands and result
            const double th_bytes = (cfg ? 4.0 : 3.0) * n * 4 + (fused ? (fu.m1 ? 3.0 : 2.0) * n * 4 : 0.0);
            if (prof_mark(h, TMD_K_SOLVER, 0, th_bytes, s, true, fused ? "x0_threshold+update" : "x0_threshold")) return 1;
            CK(x0_threshold(h->slots[ev.src], net_out, cfg ? net_out + n : nullptr, h->slots[ev.dst], ev.alpha_t, ev.sigma_t, noise_type,
                            plan->guidance_scale, Bx, (int)n_per, plan->q_lo, plan->q_weight, plan->max_val, h->d_th_ws, true, fused ? &fu : nullptr, s));
            if (prof_mark(h, TMD_K_SOLVER, 0, 0, s, false)) return 1;
            h->launches += 4;
            ++n_eval;
            if (fused) ++i;
        } else {
            const tmd_update& u = st.update;
            if (!slot_ok(u.dst) || !slot_ok(u.x) || !slot_ok(u.m) || u.m1 >= TMD_NUM_SLOTS) return fail("tmd_sample: bad slot in step %d", i);
            if (u.form < 0 || u.form > 2 || (u.form >= 1 && !slot_ok(u.m1)) || (u.form == 2 && !slot_ok(u.m2)))
                return fail("tmd_sample: bad update form / slots in step %d", i);
            const double tensors = u.form == 2 ? 5.0 : (u.m1 >= 0 ? 4.0 : 3.0);
            if (prof_mark(h, TMD_K_SOLVER, 0, tensors * n * 4, s, true, "solver_update")) return 1;
            if (u.form == 0)
                CK(solver_update(h->slots[u.dst], h->slots[u.x], h->slots[u.m], u.m1 >= 0 ? h->slots[u.m1] : nullptr, u.c1, u.c2, u.c3, n, s));
            else
                CK(solver_update_multistep(u.form, h->slots[u.dst], h->slots[u.x], h->slots[u.m], h->slots[u.m1],
                                           u.form == 2 ? h->slots[u.m2] : nullptr, u.c1, u.c2, u.c3, u.k, n, s));
            if (prof_mark(h, TMD_K_SOLVER, 0, 0, s, false)) return 1;
            ++h->launches;
        }
    }
    return 0;
}

static uint64_t fnv1a(uint64_t hsh, const void* p, size_t n) {
    const unsigned char* c = (const unsigned char*)p;
    for (size_t i = 0; i < n; ++i) { hsh ^= c[i]; hsh *= 1099511628211ull; }
    return hsh;
}

extern "C" int tmd_sample(tmd_handle h, const float* x_T, int32_t B, int32_t bands, int32_t H, int32_t W, const tmd_plan* plan, float* out,
                          float* trace, float* trace_x, void* stream) {
    if (!h || !x_T || !plan || !out || !plan->steps) return fail("tmd_sample: null argument");
    if (h->B == 0 || h->step_ops.empty()) return fail("tmd_sample: call tmd_precompute_condition first");
    cudaStream_t s = (cudaStream_t)stream;
    CK(cudaSetDevice(h->device));
    const bool cfg = plan->cfg != 0;
    if (cfg && (h->B % 2)) return fail("tmd_sample: a classifier-free-guidance plan needs 2B bound patches [uncond | cond], got %d", h->B);
    if (plan->model_type != TMD_MODEL_X_START && plan->model_type != TMD_MODEL_NOISE) return fail("tmd_sample: bad model_type %d", plan->model_type);
    const int Bx = cfg ? h->B / 2 : h->B;
    const size_t n_per = (size_t)h->N * h->H * h->W, n = n_per * Bx;
    if (B != Bx || bands != h->N || H != h->H || W != h->W)
        return fail("tmd_sample: x_T is [%d,%d,%d,%d] but the bound condition holds %d patches of [%d,%d,%d]%s", B, bands, H, W, Bx, h->N, h->H,
                    h->W, cfg ? " (classifier-free guidance: half of the 2B bound patches)" : "");
    if (n_per > (size_t)INT32_MAX) return fail("tmd_sample: patch too large");
    if (plan->q_lo < 0 || (size_t)plan->q_lo >= n_per) return fail("tmd_sample: quantile position %d outside the patch (%zu values)", plan->q_lo, n_per);
    if (plan->result_slot < 0 || plan->result_slot >= TMD_NUM_SLOTS) return fail("tmd_sample: bad result slot");

    // every evaluation's t_input -> device (the time tables are then computed by one launch inside sample_body)
    int n_evals = 0;
    for (int i = 0; i < plan->n_steps; ++i) n_evals += plan->steps[i].kind == 0;
    const bool pre_tb = n_evals > 0 && n_evals <= kMaxPlanEvals;
    if (pre_tb) {
        std::vector<float> tin;     // pageable: the async copy stages it before returning
        for (int i = 0; i < plan->n_steps; ++i)
            if (plan->steps[i].kind == 0) tin.push_back(plan->steps[i].eval.t_input);
        CK(cudaMemcpyAsync(h->d_tin_all, tin.data(), tin.size() * sizeof(float), cudaMemcpyHostToDevice, s));
    }
    CK(cudaMemcpyAsync(h->slots[0], x_T, n * sizeof(float), cudaMemcpyDeviceToDevice, s));

    static const bool graphs_on = !getenv("TMD_NO_GRAPH");
    cudaStreamCaptureStatus capturing = cudaStreamCaptureStatusNone;
    cudaStreamIsCapturing(s, &capturing);
    if (graphs_on && pre_tb && !trace && !trace_x && !h->profiling && capturing == cudaStreamCaptureStatusNone) {
        uint64_t key = fnv1a(1469598103934665603ull, plan->steps, (size_t)plan->n_steps * sizeof(tmd_step));
        const int32_t scal[6] = {plan->n_steps, plan->result_slot, plan->q_lo, plan->model_type, plan->cfg, h->B};
        const float fsc[3] = {plan->q_weight, plan->max_val, plan->guidance_scale};
        key = fnv1a(fnv1a(key, scal, sizeof(scal)), fsc, sizeof(fsc));
        tmd_engine::SampleGraph* g = nullptr;
        for (auto& c : h->graphs)
            if (c.key == key) g = &c;
        if (!g) {
            if (!h->cap_stream) CK(cudaStreamCreateWithFlags(&h->cap_stream, cudaStreamNonBlocking));
            const int64_t l0 = h->launches;
            CK(cudaStreamBeginCapture(h->cap_stream, cudaStreamCaptureModeRelaxed));
            const int rc = sample_body(h, plan, nullptr, nullptr, n_evals, pre_tb, h->cap_stream);
            cudaGraph_t graph = nullptr;
            const cudaError_t ce = cudaStreamEndCapture(h->cap_stream, &graph);
            const int64_t counted = h->launches - l0;
            h->launches = l0;
            if (rc) { if (graph) cudaGraphDestroy(graph); return 1; }
            if (ce != cudaSuccess) return fail("tmd_sample: graph capture failed: %s", cudaGetErrorString(ce));
            cudaGraphExec_t exec = nullptr;
            const cudaError_t ie = cudaGraphInstantiate(&exec, graph, 0);
            cudaGraphDestroy(graph);
            if (ie != cudaSuccess) return fail("tmd_sample: graph instantiation failed: %s", cudaGetErrorString(ie));
            if (h->graphs.size() >= 8) { cudaGraphExecDestroy(h->graphs.front().exec); h->graphs.erase(h->graphs.begin()); }
            h->graphs.push_back({key, exec, counted});
            g = &h->graphs.back();
        }
        CK(cudaGraphLaunch(g->exec, s));
        h->launches += g->launches;
    } else if (sample_body(h, plan, trace, trace_x, n_evals, pre_tb, s)) {
        return 1;
    }
    CK(cudaMemcpyAsync(out, h->slots[plan->result_slot], n * sizeof(float), cudaMemcpyDeviceToDevice, s));
    return 0;
}

extern "C" int64_t tmd_launch_count(tmd_handle h) { return h ? h->launches : 0; }

extern "C" int tmd_debug_check_redzones(tmd_handle h, int32_t* n_checked, int32_t* n_bad) {
    if (!h || !n_checked || !n_bad) return fail("tmd_debug_check_redzones: null argument");
    *n_checked = *n_bad = 0;
    CK(cudaSetDevice(h->device));
    CK(cudaDeviceSynchronize());
    std::vector<unsigned char> host(kRedzoneBytes);
    std::string first;
    for (auto& rz : h->redzones) {
        CK(cudaMemcpy(host.data(), rz.second, kRedzoneBytes, cudaMemcpyDeviceToHost));
        bool ok = true;
        for (unsigned char c : host) ok = ok && c == 0xCB;
        ++*n_checked;
        if (!ok) { ++*n_bad; if (first.empty()) first = rz.first; }
    }
    if (*n_bad) return fail("tmd_debug_check_redzones: %d of %d red zones overwritten (first: behind '%s')", *n_bad, *n_checked, first.c_str());
    return 0;
}

extern "C" int tmd_profile_begin(tmd_handle h) {
    if (!h) return fail("tmd_profile_begin: null handle");
    h->profiling = true;
    h->prof_used = 0;
    return 0;
}

extern "C" int tmd_profile_end(tmd_handle h, tmd_profile* out, const char* csv_path) {
    if (!h || !out) return fail("tmd_profile_end: null argument");
    if (!h->profiling) return fail("tmd_profile_end: tmd_profile_begin was not called");
    h->profiling = false;
    memset(out, 0, sizeof(*out));
    CK(cudaSetDevice(h->device));
    CK(cudaDeviceSynchronize());
    FILE* csv = csv_path ? fopen(csv_path, "w") : nullptr;
    if (csv) fprintf(csv, "index,kind,label,ms,flops,bytes,flops_executed\n");
    for (size_t i = 0; i < h->prof_used; ++i) {
        const ProfSpan& sp = h->prof_events[i];
        float ms = 0.f;
        CK(cudaEventElapsedTime(&ms, sp.a, sp.b));
        if (csv) fprintf(csv, "%zu,%d,%s,%.6f,%.0f,%.0f,%.0f\n", i, sp.kind, sp.label ? sp.label : "", ms, sp.flops, sp.bytes, sp.flops_exec);
        tmd_profile_kind& k = out->kind[sp.kind];
        k.ms += ms; k.flops += sp.flops; k.bytes += sp.bytes; k.launches += 1; k.flops_executed += sp.flops_exec;
    }
    if (csv) fclose(csv);
    h->prof_used = 0;
    return 0;
}

// ---------------------------------------------------------------------------------------------
// stand-alone operators
// ---------------------------------------------------------------------------------------------
extern "C" int tmd_dwt2d(const float* x, float* ll, float* lh, float* hl, float* hh, int64_t BC, int32_t H, int32_t W, void* stream) {
    if (BC == 0) return 0;
    if (!x || !ll) return fail("tmd_dwt2d: null argument");
    if ((lh || hl || hh) && !(lh && hl && hh)) return fail("tmd_dwt2d: give all three high bands or none");
    if (H % 2 || W % 2 || H <= 0 || W <= 0) return fail("tmd_dwt2d: H and W must be even (the reference mishandles odd sizes, DWT_IDWT_layer.py:1-4)");
    if (BC == 0) return 0;
    CK(dwt2d_nchw(x, ll, lh, hl, hh, BC, H, W, (cudaStream_t)stream));
    return 0;
}

extern "C" int tmd_idwt2d(const float* ll, const float* lh, const float* hl, const float* hh, float* out, int64_t BC, int32_t h, int32_t w,
                          void* stream) {
    if (BC == 0) return 0;
    if (!ll || !lh || !hl || !hh || !out) return fail("tmd_idwt2d: null argument");
    if (h <= 0 || w <= 0) return fail("tmd_idwt2d: bad size");
    if (BC == 0) return 0;
    CK(idwt2d_nchw(ll, lh, hl, hh, out, BC, h, w, (cudaStream_t)stream));
    return 0;
}

extern "C" int tmd_dynamic_threshold(float* x0, int32_t B, int32_t n, int32_t q_lo, float q_weight, float max_val, void* stream) {
    if (!x0) return fail("tmd_dynamic_threshold: null argument");
    if (B <= 0 || n <= 0 || q_lo < 0 || q_lo >= n) return fail("tmd_dynamic_threshold: bad sizes");
    cudaStream_t s = (cudaStream_t)stream;
    void* ws = nullptr;
    CK(cudaMallocAsync(&ws, threshold_workspace_bytes(B), s));
    const cudaError_t err = dynamic_threshold_inplace(x0, B, n, q_lo, q_weight, max_val, ws, s);
    cudaFreeAsync(ws, s);
    if (err != cudaSuccess) return fail("tmd_dynamic_threshold: %s", cudaGetErrorString(err));
    return 0;
}

namespace {
// The stand-alone operators take their scratch from the stream-ordered pool; keep freed blocks cached instead of
// returning them to the driver at every synchronisation (the default release threshold is 0).
void keep_pool_warm() {
    static bool done[64] = {};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64 || done[dev]) return;
    cudaMemPool_t pool;
    if (cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess) {
        uint64_t keep = ~0ull;
        cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
    }
    done[dev] = true;
}
__global__ void pack_weights_kernel(const float* __restrict__ w, float* __restrict__ pf, __nv_bfloat16* __restrict__ pb, int Co, int Ci,
                                    int taps) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t total = (size_t)Co * Ci * taps;
    if (i >= total) return;
    const int t = (int)(i % taps);
    const int ci = (int)((i / taps) % Ci);
    const int co = (int)(i / ((size_t)taps * Ci));
    const float v = w[i];
    pf[((size_t)t * Ci + ci) * Co + co] = v;
    pb[(size_t)co * taps * Ci + (size_t)t * Ci + ci] = __float2bfloat16(v);
}
}  // namespace

extern "C" int tmd_conv3d_mod(const float* x, const float* w, const float* sc, const float* bias, float* y, int32_t B, int32_t Ci,
                              int32_t Co, int32_t N, int32_t H, int32_t W, int32_t k, int32_t dtype, void* stream) {
    if (!x || !w || !y) return fail("tmd_conv3d_mod: null argument");
    if (k != 1 && k != 3) return fail("tmd_conv3d_mod: kernel size must be 1 or 3");
    if (Ci % 4 || Co % 4) return fail("tmd_conv3d_mod: Ci and Co must be multiples of 4");
    if (dtype != TMD_F32 && dtype != TMD_BF16) return fail("tmd_conv3d_mod: bad dtype");
    cudaStream_t s = (cudaStream_t)stream;
    keep_pool_warm();
    const size_t nhw = (size_t)N * H * W, es = dtype == TMD_F32 ? 4 : 2;
    const int taps = k * k * k;
    void *xin = nullptr, *yout = nullptr;
    float* pf = nullptr;
    __nv_bfloat16* pb = nullptr;
    CK(cudaMallocAsync(&xin, (size_t)B * nhw * Ci * es, s));
    CK(cudaMallocAsync(&yout, (size_t)B * nhw * Co * es, s));
    CK(cudaMallocAsync((void**)&pf, (size_t)taps * Ci * Co * 4, s));
    CK(cudaMallocAsync((void**)&pb, (size_t)taps * Ci * Co * 2, s));
    const size_t total = (size_t)Co * Ci * taps;
    pack_weights_kernel<<<(unsigned)((total + 255) / 256), 256, 0, s>>>(w, pf, pb, Co, Ci, taps);
    CK(cudaGetLastError());
    CK(ncdhw_to_cl(dtype, x, sc, xin, B, Ci, nhw, s));   // x * s == modulating the weights (no demodulation, :68-76)
    ConvOp op{};
    op.B = B; op.N = N; op.H = H; op.W = W; op.Ci = Ci; op.ld_in = Ci; op.src = xin; op.Co = Co; op.taps = taps;
    op.w_f32 = pf; op.w_bf16 = pb;
    op.epi.mode = CONV_NORMAL; op.epi.bias = bias; op.epi.raw_out = yout; op.epi.ld_raw = Co;
    int rc = 0;
    if (dtype == TMD_F32) {
        CK(conv_simt_f32(op, s));
    } else {
        // The tensor-core kernel takes up to 256 output channels per launch (one accumulator row block): wider layers -- the
        // stacked q / k / v projections of the attention modules, nn.Linear with 512+ outputs -- run as channel groups of 256,
        // each writing its slice of the channels-last output.
        const int group = Co > 256 && Co % 16 == 0 ? 256 : Co;
        ConvOp first = op;
        first.Co = group;
        if (conv_umma_supported(first)) {
            for (int co0 = 0; co0 < Co && !rc; co0 += group) {
                ConvOp g = op;
                g.Co = Co - co0 < group ? Co - co0 : group;
                g.w_bf16 = pb + (size_t)co0 * taps * Ci;
                g.epi.bias = bias ? bias + co0 : nullptr;
                g.epi.raw_out = reinterpret_cast<__nv_bfloat16*>(yout) + co0;
                if (!conv_umma_supported(g)) { rc = fail("tmd_conv3d_mod: channel group %d..%d is not supported by the tensor-core kernel", co0, co0 + g.Co); break; }
                UmmaPlan* plan = conv_umma_plan_create(g);
                if (!plan) { rc = fail("tmd_conv3d_mod: tensor map creation failed"); break; }
                const cudaError_t err = conv_umma_bf16(g, plan, s);
                conv_umma_plan_destroy(plan);
                if (err != cudaSuccess) rc = fail("tmd_conv3d_mod: %s", cudaGetErrorString(err));
            }
        } else {
            CK(conv_simt_bf16(op, s));
        }
    }
    if (!rc) CK(cl_to_ncdhw(dtype, yout, y, B, Co, nhw, s));
    cudaFreeAsync(xin, s); cudaFreeAsync(yout, s); cudaFreeAsync(pf, s); cudaFreeAsync(pb, s);
    return rc;
}

namespace {
__global__ void __launch_bounds__(256) cast_f32_bf16_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ y, size_t n8) {
    const size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (i >= n8) return;
    const float4 a = reinterpret_cast<const float4*>(x)[2 * i], b = reinterpret_cast<const float4*>(x)[2 * i + 1];
    __nv_bfloat162 o[4] = {__floats2bfloat162_rn(a.x, a.y), __floats2bfloat162_rn(a.z, a.w), __floats2bfloat162_rn(b.x, b.y),
                           __floats2bfloat162_rn(b.z, b.w)};
    reinterpret_cast<uint4*>(y)[i] = *reinterpret_cast<const uint4*>(o);
}
__global__ void __launch_bounds__(256) cast_bf16_f32_kernel(const __nv_bfloat16* __restrict__ x, float* __restrict__ y, size_t n8) {
    const size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (i >= n8) return;
    const uint4 r = reinterpret_cast<const uint4*>(x)[i];
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&r);
    const float2 a = __bfloat1622float2(h[0]), b = __bfloat1622float2(h[1]), c = __bfloat1622float2(h[2]), d = __bfloat1622float2(h[3]);
    reinterpret_cast<float4*>(y)[2 * i] = make_float4(a.x, a.y, b.x, b.y);
    reinterpret_cast<float4*>(y)[2 * i + 1] = make_float4(c.x, c.y, d.x, d.y);
}
}  // namespace

// nn.Linear / a 1x1 convolution over TOKEN-MAJOR activations: y[m, :] = W x[m, :] + bias for x [M, Ci], W [Co, Ci] (fp32),
// y [M, Co].  The rows are the "voxels" of the channels-last conv kernels, so nothing is re-laid-out: in bf16 mode x and y may
// be bf16 (x_bf16 / y_bf16) and then the call is the weight pack plus the tensor-core launches (channel groups of 256); fp32
// operands are cast on the way in / out.  fp32 mode runs the CUDA-core kernel on fp32 operands.
extern "C" int tmd_linear(const void* x, const float* w, const float* bias, void* y, int64_t M, int32_t Ci, int32_t Co, int32_t x_bf16,
                          int32_t y_bf16, int32_t dtype, void* stream) {
    if (!x || !w || !y) return fail("tmd_linear: null argument");
    if (M <= 0 || Ci <= 0 || Co <= 0 || M > 0x7fffffffLL) return fail("tmd_linear: bad sizes");
    if (Ci % 4 || Co % 4) return fail("tmd_linear: Ci and Co must be multiples of 4");
    if (dtype != TMD_F32 && dtype != TMD_BF16) return fail("tmd_linear: bad dtype");
    if (dtype == TMD_F32 && (x_bf16 || y_bf16)) return fail("tmd_linear: fp32 mode takes and returns fp32");
    cudaStream_t s = (cudaStream_t)stream;
    keep_pool_warm();
    const int Wd = M % 8 == 0 ? 8 : (M % 4 == 0 ? 4 : (M % 2 == 0 ? 2 : 1));
    float* pf = nullptr;
    __nv_bfloat16* pb = nullptr;
    void *xin = nullptr, *yout = nullptr;
    CK(cudaMallocAsync((void**)&pf, (size_t)Ci * Co * 4, s));
    CK(cudaMallocAsync((void**)&pb, (size_t)Ci * Co * 2, s));
    const size_t total = (size_t)Co * Ci;
    pack_weights_kernel<<<(unsigned)((total + 255) / 256), 256, 0, s>>>(w, pf, pb, Co, Ci, 1);
    CK(cudaGetLastError());
    ConvOp op{};
    op.B = 1; op.N = 1; op.H = (int)(M / Wd); op.W = Wd; op.Ci = Ci; op.ld_in = Ci; op.Co = Co; op.taps = 1;
    op.w_f32 = pf; op.w_bf16 = pb;
    op.epi.mode = CONV_NORMAL; op.epi.bias = bias; op.epi.ld_raw = Co;
    int rc = 0;
    if (dtype == TMD_F32) {
        op.src = x; op.epi.raw_out = y;
        const cudaError_t err = conv_simt_f32(op, s);
        if (err != cudaSuccess) rc = fail("tmd_linear: %s", cudaGetErrorString(err));
    } else {
        const size_t nx = (size_t)M * Ci, ny = (size_t)M * Co;
        if (!x_bf16) {
            if (nx % 8) rc = fail("tmd_linear: M*Ci must be a multiple of 8");
            else {
                CK(cudaMallocAsync(&xin, nx * 2, s));
                cast_f32_bf16_kernel<<<(unsigned)((nx / 8 + 255) / 256), 256, 0, s>>>((const float*)x, (__nv_bfloat16*)xin, nx / 8);
            }
        }
        if (!rc && !y_bf16) {
            if (ny % 8) rc = fail("tmd_linear: M*Co must be a multiple of 8");
            else CK(cudaMallocAsync(&yout, ny * 2, s));
        }
        if (!rc) {
            op.src = x_bf16 ? x : xin;
            __nv_bfloat16* yb = reinterpret_cast<__nv_bfloat16*>(y_bf16 ? y : yout);
            op.epi.raw_out = yb;
            const int group = Co > 256 && Co % 16 == 0 ? 256 : Co;
            ConvOp first = op;
            first.Co = group;
            if (conv_umma_supported(first)) {
                for (int co0 = 0; co0 < Co && !rc; co0 += group) {
                    ConvOp g = op;
                    g.Co = Co - co0 < group ? Co - co0 : group;
                    g.w_bf16 = pb + (size_t)co0 * Ci;
                    g.epi.bias = bias ? bias + co0 : nullptr;
                    g.epi.raw_out = yb + co0;
                    if (!conv_umma_supported(g)) { rc = fail("tmd_linear: channel group %d..%d is not supported by the tensor-core kernel", co0, co0 + g.Co); break; }
                    UmmaPlan* plan = conv_umma_plan_create(g);
                    if (!plan) { rc = fail("tmd_linear: tensor map creation failed"); break; }
                    const cudaError_t err = conv_umma_bf16(g, plan, s);
                    conv_umma_plan_destroy(plan);
                    if (err != cudaSuccess) rc = fail("tmd_linear: %s", cudaGetErrorString(err));
                }
            } else {
                const cudaError_t err = conv_simt_bf16(op, s);
                if (err != cudaSuccess) rc = fail("tmd_linear: %s", cudaGetErrorString(err));
            }
            if (!rc && !y_bf16) cast_bf16_f32_kernel<<<(unsigned)((ny / 8 + 255) / 256), 256, 0, s>>>(yb, (float*)y, ny / 8);
            if (!rc && cudaGetLastError() != cudaSuccess) rc = fail("tmd_linear: cast launch failed");
        }
    }
    if (xin) cudaFreeAsync(xin, s);
    if (yout) cudaFreeAsync(yout, s);
    cudaFreeAsync(pf, s); cudaFreeAsync(pb, s);
    return rc;
}

// ---------------------------------------------------------------------------------------------
// operators named by the north star outside the sampling path
// ---------------------------------------------------------------------------------------------
extern "C" int tmd_groupnorm(const float* x, float* y, const float* gamma, const float* beta, int32_t per_sample_affine, int32_t B,
                             int32_t C, int64_t S, int32_t G, float eps, int32_t swish, void* stream) {
    if (!x || !y) return fail("tmd_groupnorm: null argument");
    if (B <= 0 || C <= 0 || S <= 0 || G <= 0 || C % G) return fail("tmd_groupnorm: C=%d must be a positive multiple of num_groups=%d", C, G);
    cudaStream_t s = (cudaStream_t)stream;
    keep_pool_warm();
    // enough CTAs per group to fill the machine at small B*G, slices of at least 4096 values
    const size_t n = (size_t)(C / G) * (size_t)S;
    int split = (int)((4 * 148 + (size_t)B * G - 1) / ((size_t)B * G));
    const int max_split = (int)((n + 4095) / 4096);
    if (split > max_split) split = max_split;
    if (split > 64) split = 64;
    if (split < 1) split = 1;
    float* partial = nullptr;
    CK(cudaMallocAsync((void**)&partial, (size_t)B * G * split * 3 * sizeof(float), s));
    const cudaError_t err = groupnorm_affine(x, y, partial, gamma, beta, per_sample_affine ? C : 0, B, C, (size_t)S, G, split, eps, swish, s);
    cudaFreeAsync(partial, s);
    if (err != cudaSuccess) return fail("tmd_groupnorm: %s", cudaGetErrorString(err));
    return 0;
}

extern "C" int tmd_attention(const float* q, const float* k, const float* v, float* out, int32_t BH, int32_t Lq, int32_t Lk, int32_t D,
                             float scale, void* stream) {
    if (!q || !k || !v || !out) return fail("tmd_attention: null argument");
    if (BH <= 0 || Lq <= 0 || Lk <= 0 || D <= 0) return fail("tmd_attention: empty problem");
    if (D > 2048) return fail("tmd_attention: head width %d exceeds 2048", D);
    if (BH > 65535) return fail("tmd_attention: batch*heads %d exceeds 65535", BH);
    const cudaError_t err = attention_f32(q, k, v, out, BH, Lq, Lk, D, scale, (cudaStream_t)stream);
    if (err != cudaSuccess) return fail("tmd_attention: %s", cudaGetErrorString(err));
    return 0;
}

extern "C" int tmd_attention_bf16(const void* q, const void* k, const void* vt, float* out, int32_t BH, int32_t Lq, int32_t Lk,
                                  int32_t D, int32_t lk_pitch, float scale, void* stream) {
    if (!q || !k || !vt || !out) return fail("tmd_attention_bf16: null argument");
    if (!attention_bf16_supported(BH, Lq, Lk, D, lk_pitch))
        return fail("tmd_attention_bf16: unsupported problem BH=%d Lq=%d Lk=%d D=%d lk_pitch=%d (D must be a multiple of 64, lk_pitch a "
                    "multiple of 8 and >= Lk, BH <= 65535)", BH, Lq, Lk, D, lk_pitch);
    if (!(scale > 0.f)) return fail("tmd_attention_bf16: the logit scale must be positive");
    const cudaError_t err = attention_bf16(q, k, vt, out, BH, Lq, Lk, D, lk_pitch, scale, (cudaStream_t)stream);
    if (err != cudaSuccess) return fail("tmd_attention_bf16: %s", cudaGetErrorString(err));
    return 0;
}

extern "C" int tmd_attention_bf16_strided(const void* q, const void* k, const void* vt, void* out, int32_t B, int32_t heads, int32_t Lq,
                                          int32_t Lk, int32_t D, int64_t q_pitch, int64_t k_pitch, int32_t lk_pitch, int64_t out_pitch,
                                          int32_t out_bf16, float scale, void* stream) {
    if (!q || !k || !vt || !out) return fail("tmd_attention_bf16_strided: null argument");
    if (B <= 0 || heads <= 0 || (int64_t)B * heads > 65535 || !attention_bf16_supported(B * heads, Lq, Lk, D, lk_pitch))
        return fail("tmd_attention_bf16_strided: unsupported problem B=%d heads=%d Lq=%d Lk=%d D=%d lk_pitch=%d (D must be a multiple of 64, "
                    "lk_pitch a multiple of 8 and >= Lk, B*heads <= 65535)", B, heads, Lq, Lk, D, lk_pitch);
    if (q_pitch < (int64_t)heads * D || k_pitch < (int64_t)heads * D || out_pitch < (int64_t)heads * D || q_pitch % 8 || k_pitch % 8 || out_pitch % 8)
        return fail("tmd_attention_bf16_strided: row pitches must be multiples of 8 elements and hold heads*D = %d channels", heads * D);
    if (!(scale > 0.f)) return fail("tmd_attention_bf16_strided: the logit scale must be positive");
    const cudaError_t err = attention_bf16_strided(q, k, vt, out, B, heads, Lq, Lk, D, q_pitch, k_pitch, lk_pitch, out_pitch, out_bf16, scale,
                                                   (cudaStream_t)stream);
    if (err != cudaSuccess) return fail("tmd_attention_bf16_strided: %s", cudaGetErrorString(err));
    return 0;
}

extern "C" int tmd_attention_bf16_tokens(const void* q, const void* k, const void* v, void* out, int32_t B, int32_t heads, int32_t Lq,
                                         int32_t Lk, int32_t D, int64_t q_pitch, int64_t k_pitch, int64_t v_pitch, int64_t out_pitch,
                                         int32_t out_bf16, float scale, void* stream) {
    if (!q || !k || !v || !out) return fail("tmd_attention_bf16_tokens: null argument");
    if (B <= 0 || heads <= 0 || (int64_t)B * heads > 65535 || !attention_bf16_supported(B * heads, Lq, Lk, D, (Lk + 7) / 8 * 8))
        return fail("tmd_attention_bf16_tokens: unsupported problem B=%d heads=%d Lq=%d Lk=%d D=%d (D must be a multiple of 64, B*heads <= 65535)",
                    B, heads, Lq, Lk, D);
    const int64_t need = (int64_t)heads * D;
    if (q_pitch < need || k_pitch < need || v_pitch < need || out_pitch < need || q_pitch % 8 || k_pitch % 8 || v_pitch % 8 || out_pitch % 8)
        return fail("tmd_attention_bf16_tokens: row pitches must be multiples of 8 elements and hold heads*D = %d channels", heads * D);
    if (!(scale > 0.f)) return fail("tmd_attention_bf16_tokens: the logit scale must be positive");
    const cudaError_t err = attention_bf16_general(q, k, v, out, B, heads, Lq, Lk, D, q_pitch, k_pitch, v_pitch, 0, out_pitch, out_bf16, scale,
                                                   (cudaStream_t)stream);
    if (err != cudaSuccess) return fail("tmd_attention_bf16_tokens: %s", cudaGetErrorString(err));
    return 0;
}

namespace {
size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }
// scratch of one Dynamic_conv call: pooled means, expert weights, mixed weights (one layout per mode) and bias, the two layout
// conversions of the activations
size_t dynamic_conv_scratch(int B, int Ci, int Co, size_t nhw, int taps, int K, int dtype) {
    const size_t wn = (size_t)Co * Ci * taps, es = dtype == TMD_F32 ? 4 : 2;
    return align256((size_t)B * Ci * 4) + align256((size_t)B * K * 4) + align256((size_t)B * wn * 4) + align256((size_t)B * Co * 4) +
           align256((size_t)B * nhw * Ci * es) + align256((size_t)B * nhw * Co * es);
}
}  // namespace

extern "C" int64_t tmd_dynamic_conv_workspace_bytes(int32_t B, int32_t Ci, int32_t Co, int32_t N, int32_t H, int32_t W, int32_t k, int32_t K, int32_t dtype) {
    if (B <= 0 || Ci <= 0 || Co <= 0 || N <= 0 || H <= 0 || W <= 0 || (k != 1 && k != 3) || K <= 0) return -1;
    return (int64_t)dynamic_conv_scratch(B, Ci, Co, (size_t)N * H * W, k * k * k, K, dtype);
}

// One groups=B convolution (core/dynamic_conv.py:244-268) as ONE launch of the conv kernel with per-sample weight rows
// (ConvOp::w_sample_stride): spatial mean, attention head, expert mix (straight into the conv's weight layout), layout in, conv,
// layout out -- six launches, no per-sample tensor maps, one scratch block (the caller's, or one stream-ordered allocation).
static int dynamic_conv_impl(const char* who, const float* x, const float* weight, const float* bias, const float* fc1, const float* fc2,
                             const float* fc2_bias, float* y, int32_t B, int32_t Ci, int32_t Co, int32_t N, int32_t H, int32_t W, int32_t k,
                             int32_t ndim, int32_t K, int32_t hidden, float temperature, int32_t dtype, void* workspace, int64_t workspace_bytes,
                             void* stream) {
    if (!x || !weight || !fc1 || !fc2 || !y) return fail("%s: null argument", who);
    if (k != 1 && k != 3) return fail("%s: kernel size must be 1 or 3", who);
    if (Ci % 4 || Co % 4) return fail("%s: Ci and Co must be multiples of 4", who);
    if (dtype != TMD_F32 && dtype != TMD_BF16) return fail("%s: bad dtype", who);
    if (B <= 0 || K <= 0 || K > 8 || hidden <= 0) return fail("%s: need a non-empty batch, 1..8 experts and a hidden width", who);
    cudaStream_t s = (cudaStream_t)stream;
    const size_t nhw = (size_t)N * H * W, es = dtype == TMD_F32 ? 4 : 2;
    const int taps = k * k * k;
    const size_t wn = (size_t)Co * Ci * taps;
    const size_t need = dynamic_conv_scratch(B, Ci, Co, nhw, taps, K, dtype);
    char* base = (char*)workspace;
    bool own = false;
    if (!base) {
        keep_pool_warm();
        CK(cudaMallocAsync((void**)&base, need, s));
        own = true;
    } else if (workspace_bytes < (int64_t)need || (reinterpret_cast<uintptr_t>(base) & 255)) {
        return fail("%s: workspace of %lld bytes (256-byte aligned) needed, got %lld", who, (long long)need, (long long)workspace_bytes);
    }
    size_t off = 0;
    auto take = [&](size_t bytes) { char* p = base + off; off += align256(bytes); return p; };
    float* pooled = (float*)take((size_t)B * Ci * 4);
    float* alpha = (float*)take((size_t)B * K * 4);
    void* wmix = take((size_t)B * wn * 4);       // fp32 [taps][Ci][Co] for the CUDA-core kernels, bf16 [Co][taps*Ci] for the tcgen05 one
    float* ab = (float*)take((size_t)B * Co * 4);
    void* xin = take((size_t)B * nhw * Ci * es);
    void* yout = take((size_t)B * nhw * Co * es);
    int rc = 0;
    auto ck = [&](cudaError_t e, const char* what) { if (e != cudaSuccess && !rc) rc = fail("%s: %s -> %s", who, what, cudaGetErrorString(e)); };
    ck(spatial_mean(x, pooled, (size_t)B * Ci, nhw, s), "spatial_mean");
    ck(dynconv_head(pooled, fc1, fc2, fc2_bias, alpha, B, Ci, hidden, K, temperature, s), "attention head");       // attention1d / 2d carry a bias on fc2 (:18, :108)
    ConvOp op{};
    op.B = B; op.N = N; op.H = H; op.W = W; op.Ci = Ci; op.ld_in = Ci; op.Co = Co; op.taps = taps;
    op.src = xin;
    op.w_f32 = (const float*)wmix; op.w_bf16 = (const __nv_bfloat16*)wmix;
    op.w_sample_stride = (long long)wn;
    op.epi.mode = CONV_NORMAL;
    op.epi.bias = bias ? ab : nullptr; op.epi.ld_bias = bias ? Co : 0;
    op.epi.raw_out = yout; op.epi.ld_raw = Co;
    const bool tensor = dtype == TMD_BF16 && conv_umma_supported(op);        // else a CUDA-core kernel (fp32 weights)
    ck(dynconv_mix(alpha, weight, bias, tensor ? nullptr : (float*)wmix, tensor ? (__nv_bfloat16*)wmix : nullptr, ab, B, K, Co, Ci, k, ndim, s), "expert mix");
    ck(ncdhw_to_cl(dtype, x, nullptr, xin, B, Ci, nhw, s), "layout in");
    if (!rc) {
        if (dtype == TMD_F32) ck(conv_simt_f32(op, s), "conv (fp32)");
        else if (tensor) {
            UmmaPlan* plan = conv_umma_plan_create(op);
            if (!plan) rc = fail("%s: tensor map creation failed", who);
            else {
                ck(conv_umma_bf16(op, plan, s), "conv (tcgen05)");
                conv_umma_plan_destroy(plan);
            }
        } else ck(conv_simt_bf16(op, s), "conv (bf16 SIMT)");
    }
    if (!rc) ck(cl_to_ncdhw(dtype, yout, y, B, Co, nhw, s), "layout out");
    if (own) cudaFreeAsync(base, s);
    return rc;
}

extern "C" int tmd_dynamic_conv3d(const float* x, const float* weight, const float* bias, const float* fc1, const float* fc2, float* y,
                                  int32_t B, int32_t Ci, int32_t Co, int32_t N, int32_t H, int32_t W, int32_t k, int32_t K,
                                  int32_t hidden, float temperature, int32_t dtype, void* stream) {
    return dynamic_conv_impl("tmd_dynamic_conv3d", x, weight, bias, fc1, fc2, nullptr, y, B, Ci, Co, N, H, W, k, 3, K, hidden, temperature, dtype,
                             nullptr, 0, stream);
}

extern "C" int tmd_dynamic_conv(const float* x, const float* weight, const float* bias, const float* fc1, const float* fc2,
                                const float* fc2_bias, float* y, int32_t B, int32_t Ci, int32_t Co, int32_t N, int32_t H, int32_t W, int32_t k,
                                int32_t ndim, int32_t K, int32_t hidden, float temperature, int32_t dtype, void* stream) {
    if (ndim < 1 || ndim > 3) return fail("tmd_dynamic_conv: ndim must be 1, 2 or 3");
    if ((ndim < 3 && N != 1) || (ndim < 2 && H != 1)) return fail("tmd_dynamic_conv: unused leading extents must be 1");
    return dynamic_conv_impl("tmd_dynamic_conv", x, weight, bias, fc1, fc2, fc2_bias, y, B, Ci, Co, N, H, W, k, ndim, K, hidden, temperature, dtype,
                             nullptr, 0, stream);
}

extern "C" int tmd_dynamic_conv_ws(const float* x, const float* weight, const float* bias, const float* fc1, const float* fc2,
                                   const float* fc2_bias, float* y, int32_t B, int32_t Ci, int32_t Co, int32_t N, int32_t H, int32_t W, int32_t k,
                                   int32_t ndim, int32_t K, int32_t hidden, float temperature, int32_t dtype, void* workspace, int64_t workspace_bytes,
                                   void* stream) {
    if (ndim < 1 || ndim > 3) return fail("tmd_dynamic_conv_ws: ndim must be 1, 2 or 3");
    if ((ndim < 3 && N != 1) || (ndim < 2 && H != 1)) return fail("tmd_dynamic_conv_ws: unused leading extents must be 1");
    if (!workspace) return fail("tmd_dynamic_conv_ws: null workspace (size it with tmd_dynamic_conv_workspace_bytes)");
    return dynamic_conv_impl("tmd_dynamic_conv_ws", x, weight, bias, fc1, fc2, fc2_bias, y, B, Ci, Co, N, H, W, k, ndim, K, hidden, temperature, dtype,
                             workspace, workspace_bytes, stream);
}

extern "C" int tmd_quality_metrics(const float* ref, const float* img, float* out, int32_t B, int32_t C, int32_t H, int32_t W, float ratio,
                                   void* stream) {
    if (!ref || !img || !out) return fail("tmd_quality_metrics: null argument");
    if (B <= 0 || C <= 0 || C > 8 || H <= 0 || W <= 0) return fail("tmd_quality_metrics: need 1..8 bands and a non-empty image");
    const cudaError_t err = quality_metrics(ref, img, out, B, C, H * W, ratio, (cudaStream_t)stream);
    if (err != cudaSuccess) return fail("tmd_quality_metrics: %s", cudaGetErrorString(err));
    return 0;
}

extern "C" int tmd_layernorm(const float* x, float* y, const float* gamma, const float* beta, int32_t rows, int32_t C, float eps, void* stream) {
    if (!x || !y) return fail("tmd_layernorm: null argument");
    if (rows <= 0 || C <= 0) return fail("tmd_layernorm: empty problem");
    const cudaError_t err = layernorm_rows(x, y, gamma, beta, rows, C, eps, (cudaStream_t)stream);
    if (err != cudaSuccess) return fail("tmd_layernorm: %s", cudaGetErrorString(err));
    return 0;
}

extern "C" int tmd_geglu(const float* proj, float* out, int64_t rows, int32_t D, void* stream) {
    if (!proj || !out) return fail("tmd_geglu: null argument");
    if (rows <= 0 || D <= 0 || D % 4) return fail("tmd_geglu: D must be a positive multiple of 4");
    const cudaError_t err = geglu_rows(proj, out, (size_t)rows, D, (cudaStream_t)stream);
    if (err != cudaSuccess) return fail("tmd_geglu: %s", cudaGetErrorString(err));
    return 0;
}

extern "C" int tmd_band_moments(const float* a, const float* b, float* out, int32_t B, int32_t Ca, int32_t Cb, int32_t H, int32_t W,
                                int32_t highpass, void* stream) {
    if (!a || !b || !out) return fail("tmd_band_moments: null argument");
    if (B <= 0 || Ca < 1 || Cb < 1 || Ca > 8 || Cb > 8 || H <= 0 || W <= 0) return fail("tmd_band_moments: need 1..8 bands per image and a non-empty image");
    if (highpass && (H < 3 || W < 3)) return fail("tmd_band_moments: the 3x3 high-pass needs H, W >= 3");
    const cudaError_t err = band_moments(a, b, out, B, Ca, Cb, H, W, highpass != 0, (cudaStream_t)stream);
    if (err != cudaSuccess) return fail("tmd_band_moments: %s", cudaGetErrorString(err));
    return 0;
}

extern "C" int tmd_psf_downsample(const float* x, float* y, const float* param, float* psf_out, int32_t mode, int32_t B, int32_t C, int32_t H,
                                  int32_t W, int32_t k, int32_t scale, void* stream) {
    if (!x || !y || !param) return fail("tmd_psf_downsample: null argument");
    if (mode != 0 && mode != 1) return fail("tmd_psf_downsample: mode 0 (softmax PSF) or 1 (Gaussian)");
    if (B <= 0 || C <= 0 || k < 1 || k > 16 || scale < 1 || H < k || W < k) return fail("tmd_psf_downsample: need 1 <= k <= 16, scale >= 1 and an image no smaller than the kernel");
    const cudaError_t err = psf_downsample(x, y, param, psf_out, mode, B * C, H, W, k, scale, (cudaStream_t)stream);
    if (err != cudaSuccess) return fail("tmd_psf_downsample: %s", cudaGetErrorString(err));
    return 0;
}

extern "C" int tmd_scene_to_tiles(const float* pan_dn, const float* lrms_dn, float* pan_tiles, float* ms_tiles, int32_t C, int32_t Hs, int32_t Ws,
                                  int32_t ratio, int32_t tile, int32_t first, int32_t count, float img_scale, void* stream) {
    if (!pan_dn || !lrms_dn || !pan_tiles || !ms_tiles) return fail("tmd_scene_to_tiles: null argument");
    if (C <= 0 || ratio <= 0 || tile <= 0 || tile % 4 || Hs <= 0 || Ws <= 0 || Hs % tile || Ws % tile || Hs % ratio || Ws % ratio)
        return fail("tmd_scene_to_tiles: scene %dx%d must be a whole number of %d-pixel tiles (tile a multiple of 4) and of the ratio %d", Hs, Ws, tile, ratio);
    const int n_tiles = (Hs / tile) * (Ws / tile);
    if (first < 0 || count < 0 || first + count > n_tiles) return fail("tmd_scene_to_tiles: tiles [%d, %d) outside the %d tiles of the scene", first, first + count, n_tiles);
    if (!(img_scale > 0.f)) return fail("tmd_scene_to_tiles: img_scale must be positive");
    if ((reinterpret_cast<uintptr_t>(pan_dn) | reinterpret_cast<uintptr_t>(pan_tiles) | reinterpret_cast<uintptr_t>(ms_tiles)) & 15) return fail("tmd_scene_to_tiles: tensors must be 16-byte aligned");
    const cudaError_t err = scene_to_tiles(pan_dn, lrms_dn, pan_tiles, ms_tiles, C, Hs, Ws, ratio, tile, first, count, img_scale, (cudaStream_t)stream);
    if (err != cudaSuccess) return fail("tmd_scene_to_tiles: %s", cudaGetErrorString(err));
    return 0;
}

extern "C" int tmd_tiles_to_scene(const float* tiles, float* scene, int32_t C, int32_t Hs, int32_t Ws, int32_t tile, int32_t first, int32_t count,
                                  float img_scale, int32_t clamp, void* stream) {
    if (!tiles || !scene) return fail("tmd_tiles_to_scene: null argument");
    if (C <= 0 || tile <= 0 || tile % 4 || Hs <= 0 || Ws <= 0 || Hs % tile || Ws % tile)
        return fail("tmd_tiles_to_scene: scene %dx%d must be a whole number of %d-pixel tiles (tile a multiple of 4)", Hs, Ws, tile);
    const int n_tiles = (Hs / tile) * (Ws / tile);
    if (first < 0 || count < 0 || first + count > n_tiles) return fail("tmd_tiles_to_scene: tiles [%d, %d) outside the %d tiles of the scene", first, first + count, n_tiles);
    if ((reinterpret_cast<uintptr_t>(tiles) | reinterpret_cast<uintptr_t>(scene)) & 15) return fail("tmd_tiles_to_scene: tensors must be 16-byte aligned");
    const cudaError_t err = tiles_to_scene(tiles, scene, C, Hs, Ws, tile, first, count, img_scale, clamp != 0, (cudaStream_t)stream);
    if (err != cudaSuccess) return fail("tmd_tiles_to_scene: %s", cudaGetErrorString(err));
    return 0;
}

extern "C" int tmd_ancestral_update(const float* x, const float* eps, const float* noise, float* out, float c_recip, float c_recipm1,
                                    float coef1, float coef2, float sigma, int32_t clip, uint64_t seed, uint64_t step, int64_t n,
                                    void* stream) {
    if (!x || !eps || !out) return fail("tmd_ancestral_update: null argument");
    if (n <= 0 || n % 4) return fail("tmd_ancestral_update: element count must be a positive multiple of 4");
    const cudaError_t err = ancestral_update(out, x, eps, noise, c_recip, c_recipm1, coef1, coef2, sigma, clip != 0, seed, step, (size_t)n,
                                             (cudaStream_t)stream);
    if (err != cudaSuccess) return fail("tmd_ancestral_update: %s", cudaGetErrorString(err));
    return 0;
}
