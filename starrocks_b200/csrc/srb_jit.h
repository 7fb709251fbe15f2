// srb_jit.h — the TMA-staged static scan / aggregate kernel for plans that no built-in shape matches, compiled at run time.
//
// srb_shapes.cuh's kernel (k_scan_agg_static<S>) is written against a compile-time description S of the plan: expression nodes, aggregates,
// group keys. Five descriptions are compiled into libsrb.so. For every other plan the host GENERATES the description from the lowered
// program (jit_shape_of / jit_source below: a struct of constexpr tables, the same text a maintainer would have typed), hands it with the
// library's own kernel headers (embedded in the .so at build time) to NVRTC, loads the cubin through the driver API and launches it exactly
// like a built-in shape. Same kernel source, same arithmetic, same results — only the plan is no longer a list of benchmark queries.
// The reference's counterpart is the vectorised Expr / AggregateFunction tree a plan fragment instantiates per query
// (exprs/arithmetic_expr.cpp, exec/aggregate/agg_hash_variant.h: one template instantiation per key type x aggregate list, chosen at
// run time from what was compiled ahead of time); here the instantiation itself happens at run time.
//
// Compilation takes 2-4 s per distinct plan shape (NVRTC, one kernel), so it runs on a background thread and the interpreter
// (k_scan_agg_generic) serves the operator until the kernel is ready; cubins are cached per process and on disk (SRB_JIT_CACHE_DIR,
// default ~/.cache/srb_jit) under a hash of the generated source, the kernel headers and the NVRTC version. SRB_JIT = off | async (default) |
// sync; SRB_JIT_MIN_ROWS (default 4 Mi) is the batch size below which no compilation is started. srb_jit_wait() blocks until every pending
// compilation has finished (benchmarks, tests).
#pragma once
#include <dlfcn.h>
#include <sys/stat.h>
#include <unistd.h>

#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "srb_host.h"
#include "srb_shapes.cuh"

namespace srb {

// ---- the generated shape ---------------------------------------------------------------------------------------------------------
struct JitShape {
    int n_nodes = 0, n_aggs = 0, n_keys = 0, gmax = 0, threads = 0, rows = 0;
    SNode nodes[MAX_NODES];
    SAgg aggs[MAX_AGGS];
    SKey keys[MAX_KEYS];
};

// node_can_be_null / plane_kind / n_reps of srb_shapes.cuh, at run time (the geometry must be chosen before anything is compiled)
inline bool jit_node_can_be_null(const JitShape& js, int n) {
    if (n < 0) return false;
    const SNode& nd = js.nodes[n];
    if (nd.op == DOP_LOAD) return nd.nullable != 0;
    if (nd.op == DOP_LIT) return false;
    if (nd.checked) return true;
    return jit_node_can_be_null(js, nd.lhs) || jit_node_can_be_null(js, nd.rhs);
}
inline int jit_plane_kind(const JitShape& js, int a) {
    const SAgg& g = js.aggs[a];
    if (g.fn == FN_SUM || g.fn == FN_AVG) return g.dbl != 0 ? 3 : 1;
    if (g.fn == FN_COUNT) return (g.merge == 0 && g.arg >= 0 && jit_node_can_be_null(js, g.arg)) ? 1 : 0;
    return 2;
}
inline bool jit_same_plane(const JitShape& js, int a, int b) { // do two sum aggregates share an accumulator plane?
    const int ka = jit_plane_kind(js, a);
    return (ka == 1 || ka == 3) && jit_plane_kind(js, b) == ka && js.aggs[a].arg == js.aggs[b].arg && (ka == 1 || js.aggs[a].dbl == js.aggs[b].dbl);
}
inline int jit_n_reps(const JitShape& js) {
    int n = 0;
    for (int a = 0; a < js.n_aggs; a++) {
        const int kind = jit_plane_kind(js, a);
        if (kind == 0) continue;
        bool first = true;
        for (int b = 0; b < a; b++)
            if (jit_same_plane(js, a, b)) first = false;
        n += first ? 1 : 0;
    }
    return n;
}
inline bool jit_node_reads_nullable(const JitShape& js, int n) {
    if (n < 0) return false;
    const SNode& nd = js.nodes[n];
    if (nd.op == DOP_LOAD) return nd.nullable != 0;
    if (nd.op == DOP_LIT) return false;
    return jit_node_reads_nullable(js, nd.lhs) || jit_node_reads_nullable(js, nd.rhs);
}
inline int jit_n_cnt(const JitShape& js) { // planes that count the rows they took (n_cnt of srb_shapes.cuh)
    int n = 0;
    for (int a = 0; a < js.n_aggs; a++) {
        const int kind = jit_plane_kind(js, a);
        if (kind == 0 || !jit_node_reads_nullable(js, js.aggs[a].arg)) continue;
        bool first = true;
        for (int b = 0; b < a; b++)
            if (jit_same_plane(js, a, b)) first = false;
        n += first ? 1 : 0;
    }
    return n;
}
inline int jit_acc_bytes(const JitShape& js) { // AccGeom<S>::BYTES
    if (js.gmax <= 0) return 0;
    const int nr = std::max(1, jit_n_reps(js));
    return nr * js.gmax * js.threads * 8 + js.gmax * js.threads * 4 + js.threads * 8 + jit_n_cnt(js) * js.gmax * js.threads * 4;
}

// Can the static kernel run this program, and with which geometry? Integer columns and values (with or without null maps: `col_nullable`
// says which columns of THIS variant carry one), SUM / AVG / COUNT over integers, MIN / MAX over values of up to 8 bytes, up to four
// fixed-width keys that are plain columns — for ANY node / aggregate / key list.
inline bool jit_shape_of(const DProg& p, int group_hint, const bool* col_nullable, JitShape& js) {
    if (p.n_nodes < 1 || p.n_aggs < 1 || p.n_nodes > MAX_NODES) return false;
    js = JitShape{};
    js.n_nodes = p.n_nodes;
    js.n_aggs = p.n_aggs;
    js.n_keys = p.n_keys;
    int row_bytes = 0; // stage bytes per row (binary key columns: 4 bytes of offsets + 8 of window, and ~64 bytes per tile that the slack below covers)
    bool binary_col[MAX_COLS] = {};
    for (int c = 0; c < p.n_cols; c++) {
        if (p.cols[c].kind == CK_BINARY) {
            binary_col[c] = true;
            row_bytes += 12 + (col_nullable && col_nullable[c] ? 1 : 0);
            continue;
        }
        if ((p.cols[c].kind != CK_INT && p.cols[c].kind != CK_DOUBLE) || p.cols[c].width < 1 || p.cols[c].width > 16) return false;
        row_bytes += p.cols[c].width + (col_nullable && col_nullable[c] ? 1 : 0);
    }
    for (int i = 0; i < p.n_preds; i++)
        if (p.cols[p.preds[i].col].kind != CK_INT) return false; // string / DOUBLE predicates: the interpreter's
    bool col_seen[MAX_COLS] = {};
    for (int i = 0; i < p.n_nodes; i++) {
        const DNode& d = p.nodes[i];
        if (d.op < DOP_LOAD || d.op > DOP_MULFAST) return false;
        if (d.op == DOP_LOAD) {
            if (d.col < 0 || d.col >= p.n_cols || col_seen[d.col]) return false; // one LOAD per column
            if (binary_col[d.col] ? d.kind != CK_BINARY : (d.kind != p.cols[d.col].kind || p.cols[d.col].width != d.width)) return false; // (DOUBLE: an aggregate argument only)
            col_seen[d.col] = true;
        } else if (d.kind != CK_INT) {
            return false;
        } else if (d.op != DOP_LIT) {
            if (d.lhs < 0 || d.lhs >= i || (d.op != DOP_CAST && (d.rhs < 0 || d.rhs >= i))) return false;
        }
        const bool leaf = d.op == DOP_LOAD || d.op == DOP_LIT;
        if (!leaf && (p.nodes[d.lhs].kind != CK_INT || (d.op != DOP_CAST && p.nodes[d.rhs].kind != CK_INT))) return false; // strings are keys only
        js.nodes[i] = SNode{d.op, (d.op == DOP_LOAD && binary_col[d.col]) ? 0 : d.width, leaf ? -1 : d.lhs, (leaf || d.op == DOP_CAST) ? -1 : d.rhs, leaf ? 0 : d.checked,
                            (leaf || d.op == DOP_CAST) ? 0 : d.adj_l, (leaf || d.op == DOP_CAST) ? 0 : d.adj_r,
                            (d.op == DOP_LOAD && col_nullable && col_nullable[d.col]) ? 1 : 0};
    }
    for (int a = 0; a < p.n_aggs; a++) {
        const DAgg& d = p.aggs[a];
        if (d.fn < FN_SUM || d.fn > FN_MAX) return false;
        if ((d.fn == FN_MIN || d.fn == FN_MAX) && ((d.wide && !d.is_double) || d.arg < 0 || p.nodes[d.arg].width > 8)) return false;
        if (d.arg >= 0 && d.fn != FN_COUNT && (p.nodes[d.arg].kind == CK_DOUBLE) != (d.is_double != 0)) return false;
        if (d.avg_as_double && d.fn != FN_AVG) return false;
        if (d.merge && d.fn == FN_AVG) return false; // {sum, count} blobs: the interpreter's job
        if (d.fn != FN_COUNT && (d.arg < 0 || d.arg >= p.n_nodes)) return false;
        if (d.arg >= p.n_nodes) return false;
        if (d.arg >= 0 && p.nodes[d.arg].kind == CK_BINARY) return false;
        js.aggs[a] = SAgg{d.fn, d.arg, d.off_hi >= 0 ? 1 : 0, d.off_n >= 0 ? 1 : 0, d.merge ? 1 : 0, d.fn == FN_COUNT ? 0 : (d.is_double ? 1 : (d.avg_as_double ? 2 : 0))};
    }
    int key_bytes = 0;
    for (int k = 0; k < p.n_keys; k++) {
        const DKey& d = p.keys[k];
        if (p.n_keys > 4 || d.width < 0 || d.width > 8 || d.node < 0 || d.node >= p.n_nodes || p.nodes[d.node].op != DOP_LOAD) return false;
        if ((d.width == 0) != (p.nodes[d.node].kind == CK_BINARY) || p.nodes[d.node].kind == CK_DOUBLE) return false;
        js.keys[k] = SKey{d.node, d.width, d.nullable ? 1 : 0};
        key_bytes += (d.width ? d.width : 4) + (d.nullable ? 1 : 0);
    }
    if (key_bytes > 16) return false;
    for (int i = 0; i < p.n_nodes; i++) { // a binary column is loaded for one purpose: being a group key
        if (p.nodes[i].op != DOP_LOAD || !binary_col[p.nodes[i].col]) continue;
        bool is_key = false;
        for (int k = 0; k < p.n_keys; k++) is_key = is_key || p.keys[k].node == i;
        if (!is_key) return false;
    }
    for (int c = 0; c < p.n_cols; c++)
        if (binary_col[c] && !col_seen[c]) return false;
    // geometry: few groups (the planner's estimate) -> per-thread accumulators and a CTA directory of GMAX groups; otherwise every row goes to
    // the table. Threads x rows per tile: the largest tile that leaves room for >= 2 stages (3 when it can) beside the accumulators.
    const bool low = p.n_keys == 0 || (group_hint > 0 && group_hint <= 16);
    const int gmax_want = p.n_keys == 0 ? 1 : (group_hint <= 4 ? 4 : (group_hint <= 8 ? 8 : 16));
    struct Geo {
        int threads, rows;
    };
    const Geo low_geo[] = {{512, 2}, {512, 1}, {256, 2}, {256, 1}};
    const Geo flat_geo[] = {{512, 4}, {512, 2}, {512, 1}, {256, 1}}; // no GROUP BY: one accumulator set, room for a bigger tile
    const Geo high_geo[] = {{992, 2}, {992, 1}, {480, 2}, {480, 1}};
    if (const char* force = getenv("SRB_JIT_GEO")) { // experiments: "threads,rows[,gmax]"
        int t = 0, r = 0, g = -1;
        if (sscanf(force, "%d,%d,%d", &t, &r, &g) >= 2 && t > 0 && t % 32 == 0 && t <= 992 && r >= 1 && r <= 4) {
            js.threads = t;
            js.rows = r;
            js.gmax = g >= 0 ? g : (low ? gmax_want : 0);
            return jit_acc_bytes(js) + 2 * row_bytes * t * r <= ST_SMEM_BUDGET;
        }
    }
    auto pick = [&](const Geo* geo) { // the largest tile that leaves two stages beside the accumulators (measured: 512 x 2 with a ring of two beats
                                      // smaller tiles with deeper rings on every Q1 variant; the launch takes up to ST_MAX_STAGES when they fit)
        for (int i = 0; i < 4; i++) {
            js.threads = geo[i].threads;
            js.rows = geo[i].rows;
            // the end-of-kernel stash ((warp, group) x {lo, hi, nulls} per plane + the row count) is bounded by the kernel's static_assert
            if (js.gmax > 0 && (js.threads / 32) * js.gmax * (3 * std::max(1, jit_n_reps(js)) + 1) * 8 > 32 * 1024) continue;
            if (jit_acc_bytes(js) + 2 * (row_bytes * js.threads * js.rows + 64 * MAX_KEYS) <= ST_SMEM_BUDGET) return true;
        }
        return false;
    };
    if (low) {
        for (int g = gmax_want; g >= (p.n_keys == 0 ? 1 : 2); g >>= 1) { // a smaller directory before giving up the per-thread accumulators
            js.gmax = g;
            if (pick(p.n_keys == 0 ? flat_geo : low_geo)) return true;
        }
        if (p.n_keys == 0) return false;
    }
    js.gmax = 0;
    return pick(high_geo);
}

// the translation unit: the description as a struct of constexpr tables + the launch facts the host reads back
inline std::string jit_source(const JitShape& js) {
    std::string s = "#include \"srb_shapes.cuh\"\nnamespace srb {\nstruct ShapeJit {\n    static constexpr const char* name = \"jit\";\n";
    char buf[256];
    snprintf(buf, sizeof buf, "    static constexpr int N_NODES = %d, N_AGGS = %d, N_KEYS = %d, GMAX = %d, THREADS = %d, ROWS = %d;\n", js.n_nodes, js.n_aggs,
             js.n_keys, js.gmax, js.threads, js.rows);
    s += buf;
    s += "    static constexpr SNode node(int i) {\n        constexpr SNode t[N_NODES] = {\n";
    for (int i = 0; i < js.n_nodes; i++) {
        const SNode& n = js.nodes[i];
        snprintf(buf, sizeof buf, "            {%d, %d, %d, %d, %d, %d, %d, %d},\n", n.op, n.width, n.lhs, n.rhs, n.checked, n.adj_l, n.adj_r, n.nullable);
        s += buf;
    }
    s += "        };\n        return t[i];\n    }\n    static constexpr SAgg agg(int i) {\n        constexpr SAgg t[N_AGGS] = {\n";
    for (int a = 0; a < js.n_aggs; a++) {
        const SAgg& g = js.aggs[a];
        snprintf(buf, sizeof buf, "            {%d, %d, %d, %d, %d, %d},\n", g.fn, g.arg, g.wide, g.has_n, g.merge, g.dbl);
        s += buf;
    }
    s += "        };\n        return t[i];\n    }\n    static constexpr SKey key(int i) {\n";
    if (js.n_keys) {
        s += "        constexpr SKey t[N_KEYS] = {\n";
        for (int k = 0; k < js.n_keys; k++) {
            snprintf(buf, sizeof buf, "            {%d, %d, %d},\n", js.keys[k].node, js.keys[k].width, js.keys[k].nullable);
            s += buf;
        }
        s += "        };\n        return t[i];\n";
    } else {
        s += "        return SKey{0, 0};\n";
    }
    s += "    }\n};\nextern \"C\" __device__ const ShapeMeta srb_jit_meta = shape_meta<ShapeJit>();\n} // namespace srb\n";
    if (getenv("SRB_JIT_BREAK")) s += "#error SRB_JIT_BREAK is set: this translation unit is made to fail (test hook for the fallback to the interpreter)\n";
    return s;
}
constexpr const char* JIT_KERNEL_EXPR = "srb::k_scan_agg_static<srb::ShapeJit>";

} // namespace srb

// ---- NVRTC and the driver API, bound at run time (neither is needed by a process that never compiles) ----------------------------------
namespace srb {

struct NvrtcApi {
    void* h = nullptr;
    int (*CreateProgram)(void**, const char*, const char*, int, const char* const*, const char* const*) = nullptr;
    int (*DestroyProgram)(void**) = nullptr;
    int (*CompileProgram)(void*, int, const char* const*) = nullptr;
    int (*GetCUBINSize)(void*, size_t*) = nullptr;
    int (*GetCUBIN)(void*, char*) = nullptr;
    int (*GetProgramLogSize)(void*, size_t*) = nullptr;
    int (*GetProgramLog)(void*, char*) = nullptr;
    int (*AddNameExpression)(void*, const char*) = nullptr;
    int (*GetLoweredName)(void*, const char*, const char**) = nullptr;
    int (*Version)(int*, int*) = nullptr;
    std::string err;
    static NvrtcApi& get() {
        static NvrtcApi a = load();
        return a;
    }
    static NvrtcApi load() {
        NvrtcApi a;
        const char* names[] = {"libnvrtc.so.12", "libnvrtc.so", "/usr/local/cuda/lib64/libnvrtc.so.12", "/usr/local/cuda/lib64/libnvrtc.so"};
        for (const char* n : names)
            if ((a.h = dlopen(n, RTLD_NOW | RTLD_LOCAL))) break;
        if (!a.h) {
            a.err = "libnvrtc.so.12 not found";
            return a;
        }
        auto sym = [&](const char* n) {
            void* p = dlsym(a.h, n);
            if (!p && a.err.empty()) a.err = std::string("libnvrtc: missing ") + n;
            return p;
        };
        a.CreateProgram = (decltype(a.CreateProgram))sym("nvrtcCreateProgram");
        a.DestroyProgram = (decltype(a.DestroyProgram))sym("nvrtcDestroyProgram");
        a.CompileProgram = (decltype(a.CompileProgram))sym("nvrtcCompileProgram");
        a.GetCUBINSize = (decltype(a.GetCUBINSize))sym("nvrtcGetCUBINSize");
        a.GetCUBIN = (decltype(a.GetCUBIN))sym("nvrtcGetCUBIN");
        a.GetProgramLogSize = (decltype(a.GetProgramLogSize))sym("nvrtcGetProgramLogSize");
        a.GetProgramLog = (decltype(a.GetProgramLog))sym("nvrtcGetProgramLog");
        a.AddNameExpression = (decltype(a.AddNameExpression))sym("nvrtcAddNameExpression");
        a.GetLoweredName = (decltype(a.GetLoweredName))sym("nvrtcGetLoweredName");
        a.Version = (decltype(a.Version))sym("nvrtcVersion");
        return a;
    }
    bool ok() const { return h && err.empty(); }
};

struct CuDriverApi {
    void* h = nullptr;
    int (*ModuleLoadData)(void**, const void*) = nullptr;
    int (*ModuleUnload)(void*) = nullptr;
    int (*ModuleGetFunction)(void**, void*, const char*) = nullptr;
    int (*ModuleGetGlobal)(unsigned long long*, size_t*, void*, const char*) = nullptr;
    int (*FuncSetAttribute)(void*, int, int) = nullptr;
    int (*LaunchKernel)(void*, unsigned, unsigned, unsigned, unsigned, unsigned, unsigned, unsigned, void*, void**, void**) = nullptr;
    int (*GetErrorString)(int, const char**) = nullptr;
    std::string err;
    static CuDriverApi& get() {
        static CuDriverApi a = load();
        return a;
    }
    static CuDriverApi load() {
        CuDriverApi a;
        for (const char* n : {"libcuda.so.1", "libcuda.so"})
            if ((a.h = dlopen(n, RTLD_NOW | RTLD_LOCAL))) break;
        if (!a.h) {
            a.err = "libcuda.so.1 not found";
            return a;
        }
        auto sym = [&](const char* n) {
            void* p = dlsym(a.h, n);
            if (!p && a.err.empty()) a.err = std::string("libcuda: missing ") + n;
            return p;
        };
        a.ModuleLoadData = (decltype(a.ModuleLoadData))sym("cuModuleLoadData");
        a.ModuleUnload = (decltype(a.ModuleUnload))sym("cuModuleUnload");
        a.ModuleGetFunction = (decltype(a.ModuleGetFunction))sym("cuModuleGetFunction");
        a.ModuleGetGlobal = (decltype(a.ModuleGetGlobal))sym("cuModuleGetGlobal_v2");
        a.FuncSetAttribute = (decltype(a.FuncSetAttribute))sym("cuFuncSetAttribute");
        a.LaunchKernel = (decltype(a.LaunchKernel))sym("cuLaunchKernel");
        a.GetErrorString = (decltype(a.GetErrorString))sym("cuGetErrorString");
        return a;
    }
    bool ok() const { return h && err.empty(); }
    std::string msg(int rc) const {
        const char* s = nullptr;
        if (GetErrorString) GetErrorString(rc, &s);
        return s ? s : "CUDA driver error " + std::to_string(rc);
    }
};

} // namespace srb

// ---- compile, cache, load, launch ---------------------------------------------------------------------------------------------------
namespace srb {

#include "srb_jit_sources.inc" // generated by build.py: srb_jit_n_headers, srb_jit_header_names[], srb_jit_header_text[] (the kernel headers)

inline uint64_t jit_fnv(const void* p, size_t n, uint64_t h = 1469598103934665603ULL) {
    const uint8_t* b = (const uint8_t*)p;
    for (size_t i = 0; i < n; i++) h = (h ^ b[i]) * 1099511628211ULL;
    return h;
}

struct JitLoaded {
    void* module = nullptr;
    void* fn = nullptr;
    ShapeMeta meta{};
};

struct JitEntry {
    std::mutex mu;
    std::condition_variable cv;
    int state = 0; // 0: compiling, 1: cubin ready, -1: failed (err says why; the operator stays on the interpreter)
    std::string err, lowered, source;
    std::vector<char> cubin;
    double compile_ms = 0;
    bool from_disk = false;
    std::map<int, JitLoaded> loaded; // per device

    bool failed() {
        std::lock_guard<std::mutex> g(mu);
        return state == -1;
    }
    std::string error() {
        std::lock_guard<std::mutex> g(mu);
        return err;
    }
    // the kernel on `device` (the calling thread's current device), loaded on first use; nullptr while compiling or after a failure
    const JitLoaded* get(int device) {
        std::lock_guard<std::mutex> g(mu);
        if (state != 1) return nullptr;
        auto it = loaded.find(device);
        if (it != loaded.end()) return &it->second;
        CuDriverApi& drv = CuDriverApi::get();
        auto fail = [&](const std::string& why) {
            state = -1;
            err = why;
            return (const JitLoaded*)nullptr;
        };
        if (!drv.ok()) return fail(drv.err);
        cudaFree(nullptr); // the runtime's primary context is current on this thread
        JitLoaded l;
        int rc = drv.ModuleLoadData(&l.module, cubin.data());
        if (rc) return fail("cuModuleLoadData: " + drv.msg(rc));
        if ((rc = drv.ModuleGetFunction(&l.fn, l.module, lowered.c_str()))) return fail("cuModuleGetFunction: " + drv.msg(rc));
        unsigned long long dptr = 0;
        size_t bytes = 0;
        if ((rc = drv.ModuleGetGlobal(&dptr, &bytes, l.module, "srb_jit_meta")) || bytes != sizeof(ShapeMeta))
            return fail("cuModuleGetGlobal(srb_jit_meta): " + (rc ? drv.msg(rc) : std::string("size mismatch")));
        if (cudaMemcpy(&l.meta, (const void*)dptr, sizeof(ShapeMeta), cudaMemcpyDeviceToHost) != cudaSuccess) return fail("reading srb_jit_meta failed");
        if ((rc = drv.FuncSetAttribute(l.fn, 8 /* CU_FUNC_ATTRIBUTE_MAX_DYNAMIC_SHARED_SIZE_BYTES */, ST_SMEM_BUDGET)))
            return fail("cuFuncSetAttribute: " + drv.msg(rc));
        return &(loaded[device] = l);
    }
};

// NVRTC over the generated translation unit -> cubin + the kernel's mangled name
inline bool jit_compile(const std::string& source, std::vector<char>& cubin, std::string& lowered, std::string& err) {
    NvrtcApi& rtc = NvrtcApi::get();
    if (!rtc.ok()) {
        err = rtc.err;
        return false;
    }
    void* prog = nullptr;
    if (rtc.CreateProgram(&prog, source.c_str(), "srb_jit.cu", srb_jit_n_headers, srb_jit_header_text, srb_jit_header_names)) {
        err = "nvrtcCreateProgram failed";
        return false;
    }
    rtc.AddNameExpression(prog, JIT_KERNEL_EXPR);
    const char* opts[] = {"--gpu-architecture=sm_100a", "-std=c++17", "-lineinfo", "--device-int128", "-default-device"};
    const int rc = rtc.CompileProgram(prog, 5, opts);
    bool ok = rc == 0;
    if (!ok) {
        size_t n = 0;
        rtc.GetProgramLogSize(prog, &n);
        std::string log(n, '\0');
        if (n) rtc.GetProgramLog(prog, &log[0]);
        err = "nvrtcCompileProgram failed (" + std::to_string(rc) + "): " + log.substr(0, 2000);
    } else {
        const char* nm = nullptr;
        size_t n = 0;
        if (rtc.GetLoweredName(prog, JIT_KERNEL_EXPR, &nm) || !nm || rtc.GetCUBINSize(prog, &n) || n == 0) {
            err = "nvrtc: no cubin / lowered name";
            ok = false;
        } else {
            lowered = nm;
            cubin.resize(n);
            ok = rtc.GetCUBIN(prog, cubin.data()) == 0;
            if (!ok) err = "nvrtcGetCUBIN failed";
        }
    }
    rtc.DestroyProgram(&prog);
    return ok;
}

class JitCache {
public:
    static JitCache& get() {
        static JitCache* c = new JitCache(); // never destroyed: a compile thread may outlive main()
        return *c;
    }
    // the entry of this generated source; starts its compilation (inline when `sync`) the first time it is asked for
    std::shared_ptr<JitEntry> acquire(const std::string& source, bool sync) {
        std::shared_ptr<JitEntry> e;
        bool created = false;
        {
            std::lock_guard<std::mutex> g(mu_);
            auto it = entries_.find(source);
            if (it != entries_.end()) {
                e = it->second;
            } else {
                e = std::make_shared<JitEntry>();
                e->source = source;
                entries_[source] = e;
                pending_++;
                created = true;
            }
        }
        if (created && sync)
            build(e);
        else if (created)
            std::thread([this, e] { build(e); }).detach();
        else if (sync)
            wait(e); // another operator's compilation of the same shape
        return e;
    }
    std::shared_ptr<JitEntry> peek(const std::string& source) {
        std::lock_guard<std::mutex> g(mu_);
        auto it = entries_.find(source);
        return it == entries_.end() ? nullptr : it->second;
    }
    static void wait(const std::shared_ptr<JitEntry>& e) {
        std::unique_lock<std::mutex> l(e->mu);
        e->cv.wait(l, [&] { return e->state != 0; });
    }
    void wait_all() {
        std::unique_lock<std::mutex> l(mu_);
        cv_.wait(l, [&] { return pending_ == 0; });
    }
    // {compiled, loaded from disk, failed, pending}
    void stats(int64_t out[4]) {
        std::lock_guard<std::mutex> g(mu_);
        out[0] = out[1] = out[2] = 0;
        out[3] = pending_;
        for (auto& kv : entries_) {
            std::lock_guard<std::mutex> ge(kv.second->mu);
            if (kv.second->state == 1) out[kv.second->from_disk ? 1 : 0]++;
            if (kv.second->state == -1) out[2]++;
        }
    }
    std::string last_error() {
        std::lock_guard<std::mutex> g(mu_);
        for (auto& kv : entries_) {
            std::lock_guard<std::mutex> ge(kv.second->mu);
            if (kv.second->state == -1) return kv.second->err;
        }
        return "";
    }

private:
    std::mutex mu_;
    std::condition_variable cv_;
    std::map<std::string, std::shared_ptr<JitEntry>> entries_;
    int pending_ = 0;

    static std::string cache_dir() {
        const char* d = getenv("SRB_JIT_CACHE_DIR");
        if (d && *d) return d;
        const char* home = getenv("HOME");
        return std::string(home && *home ? home : "/tmp") + "/.cache/srb_jit";
    }
    static std::string cache_path(const std::string& source) {
        uint64_t h = jit_fnv(source.data(), source.size());
        for (int i = 0; i < srb_jit_n_headers; i++) h = jit_fnv(srb_jit_header_text[i], strlen(srb_jit_header_text[i]), h);
        int major = 0, minor = 0;
        if (NvrtcApi::get().ok()) NvrtcApi::get().Version(&major, &minor);
        h = jit_fnv(&major, sizeof major, jit_fnv(&minor, sizeof minor, h));
        char name[64];
        snprintf(name, sizeof name, "/%016llx.sm_100a.cubin", (unsigned long long)h);
        return cache_dir() + name;
    }
    static bool read_disk(const std::string& path, std::vector<char>& cubin, std::string& lowered) {
        FILE* f = fopen(path.c_str(), "rb");
        if (!f) return false;
        uint32_t hdr[3] = {0, 0, 0}; // magic, name length, cubin length
        bool ok = fread(hdr, 4, 3, f) == 3 && hdr[0] == 0x4a425253u && hdr[1] < 4096 && hdr[2] > 0 && hdr[2] < (64u << 20);
        if (ok) {
            lowered.resize(hdr[1]);
            cubin.resize(hdr[2]);
            ok = fread(&lowered[0], 1, hdr[1], f) == hdr[1] && fread(cubin.data(), 1, hdr[2], f) == hdr[2];
        }
        fclose(f);
        return ok;
    }
    static void write_disk(const std::string& path, const std::vector<char>& cubin, const std::string& lowered) {
        const std::string dir = cache_dir();
        std::string cur;
        for (size_t i = 1; i <= dir.size(); i++) // mkdir -p
            if (i == dir.size() || dir[i] == '/') {
                cur = dir.substr(0, i);
                mkdir(cur.c_str(), 0755);
            }
        const std::string tmp = path + "." + std::to_string((long)getpid()) + ".tmp";
        FILE* f = fopen(tmp.c_str(), "wb");
        if (!f) return;
        const uint32_t hdr[3] = {0x4a425253u, (uint32_t)lowered.size(), (uint32_t)cubin.size()};
        const bool ok = fwrite(hdr, 4, 3, f) == 3 && fwrite(lowered.data(), 1, lowered.size(), f) == lowered.size() &&
                        fwrite(cubin.data(), 1, cubin.size(), f) == cubin.size();
        fclose(f);
        if (ok)
            rename(tmp.c_str(), path.c_str()); // atomic: readers see a whole file or none
        else
            unlink(tmp.c_str());
    }
    void build(const std::shared_ptr<JitEntry>& e) {
        const auto t0 = std::chrono::steady_clock::now();
        std::vector<char> cubin;
        std::string lowered, err;
        const bool use_disk = !getenv("SRB_JIT_NO_DISK_CACHE");
        const std::string path = cache_path(e->source);
        bool from_disk = use_disk && read_disk(path, cubin, lowered);
        bool ok = from_disk || jit_compile(e->source, cubin, lowered, err);
        if (ok && !from_disk && use_disk) write_disk(path, cubin, lowered);
        {
            std::lock_guard<std::mutex> g(e->mu);
            e->cubin.swap(cubin);
            e->lowered = lowered;
            e->err = err;
            e->from_disk = from_disk;
            e->compile_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
            e->state = ok ? 1 : -1;
        }
        e->cv.notify_all();
        {
            std::lock_guard<std::mutex> g(mu_);
            pending_--;
        }
        cv_.notify_all();
    }
};

// launches the compiled kernel over the leading whole tiles of [base, base+n) like launch_shape<S> does for a built-in shape
inline int64_t jit_launch(const JitLoaded& k, const DProg& p, const TableView& tv, int64_t base, int64_t n, uint32_t* defer_out,
                          unsigned long long* rows_selected, unsigned long long* deferred, unsigned long long* inserted, int n_sms, cudaStream_t stream,
                          bool allow_narrow) {
    StaticArgs sa{};
    size_t smem = 0;
    int64_t n_tiles = prepare_static(k.meta, p, base, n, allow_narrow, sa, smem);
    if (n_tiles == 0) return 0;
    const unsigned grid = (unsigned)(n_tiles < n_sms ? n_tiles : n_sms);
    StaticCtr ctr{rows_selected, deferred, inserted};
    void* params[] = {(void*)&p, &sa, (void*)&tv, &base, &n_tiles, &defer_out, &ctr};
    CuDriverApi& drv = CuDriverApi::get();
    const int rc = drv.LaunchKernel(k.fn, grid, 1, 1, (unsigned)k.meta.threads + 32, 1, 1, (unsigned)smem, (void*)stream, params, nullptr);
    if (rc) throw Error(SRB_INTERNAL_ERROR, "cuLaunchKernel (run-time compiled shape): " + drv.msg(rc));
    return n_tiles * k.meta.tile;
}

} // namespace srb
