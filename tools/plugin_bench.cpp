// plugin_bench.cpp — times (and checks) the drop-in seat itself: one decoded token's MUL_MAT / ADD node list of a Llama-shaped
// model through the REFERENCE's public API on libggml-b200.so: ggml_backend_load -> ggml_backend_tensor_set(x) ->
// ggml_backend_graph_compute -> ggml_backend_tensor_get(logits), host buffers in and out every step.
// Test / bench infrastructure: built by oracle/Makefile.ref against the reference's headers and libggml (oracle/_ref), never part
// of the product.  Reference caller being imitated: ggml_backend_graph_compute, ggml/src/ggml-backend.cpp:327-335.
//
// usage: b200-plugin-bench <plugin.so> <llama2-7b|llama3-8b|tiny> <steps> <warmup> [--check] [--cols N]
//   --check : also run the same graph on the reference CPU backend and report the NMSE of the logits (full size)
#include "ggml.h"
#include "ggml-alloc.h"
#include "ggml-backend.h"

#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <string>
#include <vector>

struct Shape { const char * name; int64_t M, K; };
struct Model { int64_t h, f, kv, L, V; };

static Model model_of(const std::string & m) {
    if (m == "llama2-7b") return {4096, 11008, 4096, 32, 32000};
    if (m == "llama3-8b") return {4096, 14336, 1024, 32, 128256};
    if (m == "tiny")      return {1024, 2048, 256, 2, 4096};
    fprintf(stderr, "unknown model %s\n", m.c_str()); exit(2);
}

// random-init q4_0 blocks in the wire format (random nibbles, fp16 scale +-U(2e-4, 4e-4)): same recipe as bench.py
static std::vector<uint8_t> random_q4_0(int64_t rows, int64_t K, std::mt19937 & rng) {
    const int64_t nb = K / 32;
    std::vector<uint8_t> raw((size_t) rows * nb * 18);
    uint32_t * w = (uint32_t *) raw.data();
    for (size_t i = 0; i < raw.size() / 4; i++) w[i] = rng();
    for (size_t i = raw.size() / 4 * 4; i < raw.size(); i++) raw[i] = (uint8_t) rng();
    std::uniform_real_distribution<float> d(0.0002f, 0.0004f);
    for (int64_t b = 0; b < rows * nb; b++) {
        const float v = d(rng) * ((rng() & 1) ? 1.f : -1.f);
        const ggml_fp16_t h = ggml_fp32_to_fp16(v);
        memcpy(raw.data() + b * 18, &h, 2);
    }
    return raw;
}

struct Net {
    ggml_context * ctx = nullptr; ggml_backend_buffer_t buf = nullptr; ggml_cgraph * gf = nullptr;
    ggml_tensor * x = nullptr, * logits = nullptr;
    std::vector<ggml_tensor *> weights;      // order: per layer wq wk wv wo gate up down, then output
    int n_mulmat = 0, n_add = 0;
};

static Net build(const Model & md, ggml_backend_t be, int64_t N) {
    Net n;
    ggml_init_params ip = { ggml_tensor_overhead() * (size_t) (md.L * 20 + 16) + ggml_graph_overhead_custom(md.L * 12 + 16, false) + (1 << 20), nullptr, true };
    n.ctx = ggml_init(ip);
    n.gf = ggml_new_graph_custom(n.ctx, md.L * 12 + 16, false);
    n.x = ggml_new_tensor_2d(n.ctx, GGML_TYPE_F32, md.h, N);
    ggml_tensor * cur = n.x;
    auto W = [&](int64_t M, int64_t K) { ggml_tensor * t = ggml_new_tensor_2d(n.ctx, GGML_TYPE_Q4_0, K, M); n.weights.push_back(t); return t; };
    auto add_node = [&](ggml_tensor * t) { ggml_build_forward_expand(n.gf, t); return t; };
    for (int64_t l = 0; l < md.L; l++) {
        ggml_tensor * q = add_node(ggml_mul_mat(n.ctx, W(md.h, md.h), cur));
        add_node(ggml_mul_mat(n.ctx, W(md.kv, md.h), cur));
        add_node(ggml_mul_mat(n.ctx, W(md.kv, md.h), cur));
        ggml_tensor * o = add_node(ggml_mul_mat(n.ctx, W(md.h, md.h), q));        // (attention itself is not on this path: q stands in)
        ggml_tensor * h = add_node(ggml_add(n.ctx, cur, o));
        ggml_tensor * g = add_node(ggml_mul_mat(n.ctx, W(md.f, md.h), h));
        add_node(ggml_mul_mat(n.ctx, W(md.f, md.h), h));
        ggml_tensor * dn = add_node(ggml_mul_mat(n.ctx, W(md.h, md.f), g));       // (gate stands in for silu(gate) * up)
        cur = add_node(ggml_add(n.ctx, h, dn));
        n.n_mulmat += 7; n.n_add += 2;
    }
    n.logits = add_node(ggml_mul_mat(n.ctx, W(md.V, md.h), cur));
    n.n_mulmat++;
    n.buf = ggml_backend_alloc_ctx_tensors(n.ctx, be);
    if (!n.buf) { fprintf(stderr, "alloc_ctx_tensors failed\n"); exit(3); }
    return n;
}

int main(int argc, char ** argv) {
    if (argc < 5) { fprintf(stderr, "usage: %s <plugin.so> <model> <steps> <warmup> [--check] [--cols N]\n", argv[0]); return 2; }
    const std::string plugin = argv[1], model = argv[2];
    const int steps = atoi(argv[3]), warmup = atoi(argv[4]);
    bool check = false; int64_t N = 1;
    for (int i = 5; i < argc; i++) { if (!strcmp(argv[i], "--check")) check = true; else if (!strcmp(argv[i], "--cols") && i + 1 < argc) N = atoll(argv[++i]); }
    ggml_backend_reg_t reg = ggml_backend_load(plugin.c_str());
    if (!reg || ggml_backend_reg_dev_count(reg) == 0) { fprintf(stderr, "cannot load %s (or it has no device)\n", plugin.c_str()); return 3; }
    ggml_backend_dev_t dev = ggml_backend_reg_dev_get(reg, 0);
    ggml_backend_t be = ggml_backend_dev_init(dev, nullptr);
    if (!be) return 3;
    const Model md = model_of(model);
    Net net = build(md, be, N);
    // one random host matrix per shape, uploaded into every layer's own tensor (distinct HBM copies, no L2 reuse between layers)
    std::mt19937 rng(1234);
    const Shape shapes[8] = { {"wq", md.h, md.h}, {"wk", md.kv, md.h}, {"wv", md.kv, md.h}, {"wo", md.h, md.h}, {"gate", md.f, md.h}, {"up", md.f, md.h},
                              {"down", md.h, md.f}, {"output", md.V, md.h} };
    std::vector<std::vector<uint8_t>> host(8);
    for (int s = 0; s < 8; s++) host[s] = random_q4_0(shapes[s].M, shapes[s].K, rng);
    size_t wbytes = 0;
    for (size_t i = 0; i < net.weights.size(); i++) {
        const int s = i + 1 == net.weights.size() ? 7 : (int) (i % 7);
        ggml_backend_tensor_set(net.weights[i], host[s].data(), 0, host[s].size());
        wbytes += host[s].size();
    }
    std::vector<float> x((size_t) md.h * N), logits((size_t) md.V * N);
    { std::uniform_real_distribution<float> d(-1.f, 1.f); for (auto & v : x) v = d(rng); }
    auto step = [&]() {
        ggml_backend_tensor_set(net.x, x.data(), 0, x.size() * 4);
        if (ggml_backend_graph_compute(be, net.gf) != GGML_STATUS_SUCCESS) { fprintf(stderr, "graph_compute failed\n"); exit(4); }
        ggml_backend_tensor_get(net.logits, logits.data(), 0, logits.size() * 4);
    };
    for (int i = 0; i < (warmup < 1 ? 1 : warmup); i++) step();
    const auto t0 = std::chrono::steady_clock::now();
    for (int i = 0; i < steps; i++) step();
    const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() / (steps > 0 ? steps : 1);
    // the same with the stream-ordered calls a scheduler uses (set_tensor_async / graph_compute_async / get_tensor_async + synchronize)
    auto step_async = [&]() {
        ggml_backend_tensor_set_async(be, net.x, x.data(), 0, x.size() * 4);
        if (ggml_backend_graph_compute_async(be, net.gf) != GGML_STATUS_SUCCESS) exit(4);
        ggml_backend_tensor_get_async(be, net.logits, logits.data(), 0, logits.size() * 4);
        ggml_backend_synchronize(be);
    };
    for (int i = 0; i < 3; i++) step_async();
    const auto t1 = std::chrono::steady_clock::now();
    for (int i = 0; i < steps; i++) step_async();
    const double sec_async = std::chrono::duration<double>(std::chrono::steady_clock::now() - t1).count() / (steps > 0 ? steps : 1);
    bool finite = true; for (float v : logits) finite = finite && std::isfinite(v);
    double nmse = -1.0;
    if (check) {
        ggml_backend_t cpu = ggml_backend_init_by_type(GGML_BACKEND_DEVICE_TYPE_CPU, nullptr);
        Net ref = build(md, cpu, N);
        for (size_t i = 0; i < ref.weights.size(); i++) {
            const int s = i + 1 == ref.weights.size() ? 7 : (int) (i % 7);
            ggml_backend_tensor_set(ref.weights[i], host[s].data(), 0, host[s].size());
        }
        std::vector<float> want(logits.size());
        ggml_backend_tensor_set(ref.x, x.data(), 0, x.size() * 4);
        if (ggml_backend_graph_compute(cpu, ref.gf) != GGML_STATUS_SUCCESS) return 5;
        ggml_backend_tensor_get(ref.logits, want.data(), 0, want.size() * 4);
        double num = 0, den = 0;
        for (size_t i = 0; i < want.size(); i++) { const double d = (double) logits[i] - want[i]; num += d * d; den += (double) want[i] * want[i]; }
        nmse = num / (den > 0 ? den : 1e-30);
        ggml_backend_buffer_free(ref.buf); ggml_free(ref.ctx); ggml_backend_free(cpu);
    }
    char nm[32]; if (nmse < 0) snprintf(nm, sizeof(nm), "null"); else snprintf(nm, sizeof(nm), "%.3e", nmse);
    printf("{\"impl\": \"plugin\", \"backend\": \"%s\", \"model\": \"%s\", \"cols\": %lld, \"nodes\": %d, \"mul_mat\": %d, \"add\": %d, \"weight_bytes\": %zu, "
           "\"tokens_per_s\": %.3f, \"ms_per_step\": %.5f, \"tokens_per_s_async_api\": %.3f, \"h2d_bytes_per_step\": %zu, \"d2h_bytes_per_step\": %zu, "
           "\"logits_finite\": %s, \"nmse_vs_ggml_cpu\": %s}\n",
           ggml_backend_name(be), model.c_str(), (long long) N, ggml_graph_n_nodes(net.gf), net.n_mulmat, net.n_add, wbytes,
           (double) N / sec, sec * 1e3, (double) N / sec_async, x.size() * 4, logits.size() * 4, finite ? "true" : "false", nm);
    ggml_backend_buffer_free(net.buf); ggml_free(net.ctx); ggml_backend_free(be);
    return finite && (nmse < 0 || nmse <= 5e-4) ? 0 : 6;
}
