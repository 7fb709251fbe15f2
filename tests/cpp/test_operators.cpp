// tests/cpp/test_operators.cpp — drives the C++ operator binding (include/srb/gpu_operators.h) the way PipelineDriver::process
// drives the reference's operators (be/src/exec/pipeline/pipeline_driver.cpp:286-373): 4096-row chunks pushed while need_input,
// set_finishing when the source is drained, chunks of <= chunk_size rows pulled while has_output. Expected values are computed
// here with plain loops over the same rows (the reference's semantics: NULL never passes a compare, decimal product of two
// DECIMAL64 casts in int128, SUM skips NULLs, count(*) counts rows, inner join emits every matching pair).
//
//   g++ -std=c++17 -I include tests/cpp/test_operators.cpp -L starrocks_b200 -lsrb -Wl,-rpath,$PWD/starrocks_b200 -o build/test_operators
//   ./build/test_operators            (needs a B200; exits 0 and prints PASS lines)
#include <cstdio>
#include <cstdlib>
#include <map>
#include <set>
#include <tuple>

#include "srb/gpu_operators.h"

using namespace starrocks;
using namespace starrocks::pipeline;

static int g_failed = 0, g_checks = 0;
#define CHECK(cond)                                                            \
    do {                                                                       \
        g_checks++;                                                            \
        if (!(cond)) {                                                         \
            g_failed++;                                                        \
            std::printf("FAIL %s:%d: %s\n", __FILE__, __LINE__, #cond);        \
        }                                                                      \
    } while (0)

static uint64_t splitmix(uint64_t x) {
    x += 0x9e3779b97f4a7c15ULL;
    x = (x ^ (x >> 30)) * 0xbf58476d1ce4e5b9ULL;
    x = (x ^ (x >> 27)) * 0x94d049bb133111ebULL;
    return x ^ (x >> 31);
}

static srb_expr_node slot_ref(int slot, int type, int precision = 0, int scale = 0) {
    srb_expr_node n{};
    n.op = SRB_EXPR_SLOT_REF;
    n.type = type;
    n.precision = precision;
    n.scale = scale;
    n.lhs = n.rhs = -1;
    n.slot_id = slot;
    return n;
}
static srb_expr_node unary(int op, int type, int precision, int scale, int lhs) {
    srb_expr_node n{};
    n.op = op;
    n.type = type;
    n.precision = precision;
    n.scale = scale;
    n.lhs = lhs;
    n.rhs = -1;
    return n;
}
static srb_expr_node binary(int op, int type, int precision, int scale, int lhs, int rhs) {
    srb_expr_node n = unary(op, type, precision, scale, lhs);
    n.rhs = rhs;
    return n;
}
static srb_predicate pred(int slot, int op, int type, int64_t lit) {
    srb_predicate p{};
    p.slot_id = slot;
    p.op = op;
    p.type = type;
    p.lit_lo = lit;
    p.lit_hi = lit < 0 ? -1 : 0;
    return p;
}

// ---- Q6 shape: range filter + sum(cast(ext) * cast(disc)) --------------------------------------------------------------------
static void test_filter_and_decimal_sum(srb_ctx* ctx) {
    const int SHIP = 1, DISC = 2, QTY = 3, EXT = 4;
    const int64_t n = 50000;
    std::vector<ChunkPtr> chunks;
    int128_t want = 0;
    int64_t want_rows = 0;
    for (int64_t base = 0; base < n; base += 4096) {
        const int64_t m = std::min<int64_t>(4096, n - base);
        auto ship = std::make_shared<Int32Column>(SRB_TYPE_DATE);
        auto disc = std::make_shared<DecimalV3Column<int64_t>>(15, 2);
        auto qty = std::make_shared<DecimalV3Column<int64_t>>(15, 2);
        auto ext = std::make_shared<DecimalV3Column<int64_t>>(15, 2);
        for (int64_t i = base; i < base + m; i++) {
            const int32_t s = 2448623 + (int32_t)(splitmix(i) % 2500);
            const int64_t d = (int64_t)(splitmix(i + 1000000) % 11), q = 100 * (1 + (int64_t)(splitmix(i + 2000000) % 50));
            const int64_t e = 90100 + (int64_t)(splitmix(i + 3000000) % 10404850);
            ship->append(s);
            disc->append(d);
            qty->append(q);
            ext->append(e);
            if (s >= 2449354 && s < 2449719 && d >= 5 && d <= 7 && q < 2400) {
                want += (int128_t)e * (int128_t)d;
                want_rows++;
            }
        }
        auto c = std::make_shared<Chunk>();
        c->append_column(ship, SHIP);
        c->append_column(disc, DISC);
        c->append_column(qty, QTY);
        c->append_column(ext, EXT);
        chunks.push_back(c);
    }
    ScanAggPlan plan;
    plan.preds = {pred(SHIP, SRB_PRED_GE, SRB_TYPE_DATE, 2449354), pred(SHIP, SRB_PRED_LT, SRB_TYPE_DATE, 2449719),
                  pred(DISC, SRB_PRED_GE, SRB_TYPE_DECIMAL64, 5), pred(DISC, SRB_PRED_LE, SRB_TYPE_DECIMAL64, 7),
                  pred(QTY, SRB_PRED_LT, SRB_TYPE_DECIMAL64, 2400)};
    plan.nodes = {slot_ref(EXT, SRB_TYPE_DECIMAL64, 15, 2), slot_ref(DISC, SRB_TYPE_DECIMAL64, 15, 2),
                  unary(SRB_EXPR_CAST, SRB_TYPE_DECIMAL128, 15, 2, 0), unary(SRB_EXPR_CAST, SRB_TYPE_DECIMAL128, 15, 2, 1),
                  binary(SRB_EXPR_MUL, SRB_TYPE_DECIMAL128, 30, 4, 2, 3)};
    plan.aggs = {srb_agg_fn{SRB_AGG_SUM, 4}};
    RuntimeState state(4096);
    ChunkSourceOperator src(chunks);
    GpuScanAggregateOperator agg(ctx, plan);
    auto res = run_pipeline(&state, {&src, &agg});
    CHECK(res.ok());
    if (!res.ok()) {
        std::printf("  %s\n", res.status().to_string().c_str());
        return;
    }
    CHECK(res.value().size() == 1 && res.value()[0]->num_rows() == 1);
    auto col = res.value()[0]->get_column_by_index(0);
    CHECK(col->logical_type() == SRB_TYPE_DECIMAL128 && col->scale() == 4);
    const Column* data = col->is_nullable() ? static_cast<const NullableColumn*>(col.get())->data_column().get() : col.get();
    CHECK(*(const int128_t*)data->raw_data() == want);
    CHECK(want_rows > 100);
    std::printf("PASS filter + decimal sum (%lld of %lld rows selected)\n", (long long)want_rows, (long long)n);
}

// ---- GROUP BY a nullable INT key: sum(v), count(*) ; the NULL key is its own group ------------------------------------------------
static void test_group_by_nullable_key(srb_ctx* ctx) {
    const int KEY = 10, VAL = 11;
    const int64_t n = 30000;
    std::vector<ChunkPtr> chunks;
    std::map<int64_t, std::pair<int64_t, int64_t>> want; // key (INT64_MIN = NULL) -> (sum, count)
    for (int64_t base = 0; base < n; base += 4096) {
        const int64_t m = std::min<int64_t>(4096, n - base);
        auto k = std::make_shared<Int32Column>(SRB_TYPE_INT);
        auto nul = std::make_shared<NullColumn>(SRB_TYPE_BOOLEAN);
        auto v = std::make_shared<Int64Column>(SRB_TYPE_BIGINT);
        for (int64_t i = base; i < base + m; i++) {
            const int32_t key = (int32_t)(splitmix(i) % 9000) - 4500;
            const bool is_null = splitmix(i + 77) % 50 == 0;
            const int64_t val = (int64_t)(splitmix(i + 555) % 2000001) - 1000000;
            k->append(key);
            nul->append(is_null ? 1 : 0);
            v->append(val);
            auto& g = want[is_null ? INT64_MIN : key];
            g.first += val;
            g.second += 1;
        }
        auto c = std::make_shared<Chunk>();
        c->append_column(std::make_shared<NullableColumn>(k, nul), KEY);
        c->append_column(v, VAL);
        chunks.push_back(c);
    }
    ScanAggPlan plan;
    plan.nodes = {slot_ref(KEY, SRB_TYPE_INT), slot_ref(VAL, SRB_TYPE_BIGINT)};
    plan.key_nodes = {0};
    plan.aggs = {srb_agg_fn{SRB_AGG_SUM, 1}, srb_agg_fn{SRB_AGG_COUNT, -1}};
    RuntimeState state(4096);
    ChunkSourceOperator src(chunks);
    GpuScanAggregateOperator agg(ctx, plan);
    auto res = run_pipeline(&state, {&src, &agg});
    CHECK(res.ok());
    if (!res.ok()) {
        std::printf("  %s\n", res.status().to_string().c_str());
        return;
    }
    std::map<int64_t, std::pair<int64_t, int64_t>> got;
    size_t rows = 0;
    for (auto& ch : res.value()) {
        CHECK(ch->num_rows() <= 4096 && ch->num_columns() == 3);
        auto kc = ch->get_column_by_index(0);
        CHECK(kc->is_nullable());
        auto* kn = static_cast<const NullableColumn*>(kc.get());
        const int32_t* kd = (const int32_t*)kn->raw_data();
        auto sc = ch->get_column_by_index(1);
        const Column* sd = sc->is_nullable() ? static_cast<const NullableColumn*>(sc.get())->data_column().get() : sc.get();
        const int64_t* sums = (const int64_t*)sd->raw_data();
        const int64_t* cnts = (const int64_t*)ch->get_column_by_index(2)->raw_data();
        for (size_t r = 0; r < ch->num_rows(); r++, rows++) got[kn->is_null(r) ? INT64_MIN : kd[r]] = {sums[r], cnts[r]};
    }
    CHECK(rows == want.size()); // no duplicate groups
    CHECK(got == want);
    CHECK(want.count(INT64_MIN) == 1);
    std::printf("PASS group by nullable key (%zu groups in %zu chunks)\n", want.size(), res.value().size());
}

// ---- inner hash join: build (key, payload) |><| probe (key, x) ---------------------------------------------------------------------
static void test_inner_join(srb_ctx* ctx) {
    const int BK = 20, BP = 21, PK = 30, PX = 31;
    std::multimap<int32_t, int64_t> table;
    std::vector<ChunkPtr> builds; // the build side arrives in chunks of <= chunk_size rows too
    for (int base = 0; base < 5000; base += 4096) {
        auto bk = std::make_shared<Int32Column>(SRB_TYPE_INT);
        auto bp = std::make_shared<Int64Column>(SRB_TYPE_BIGINT);
        for (int i = base; i < std::min(base + 4096, 5000); i++) {
            const int32_t k = (int32_t)(splitmix(i) % 3000);
            bk->append(k);
            bp->append(1000 + i);
            table.emplace(k, 1000 + i);
        }
        auto build = std::make_shared<Chunk>();
        build->append_column(bk, BK);
        build->append_column(bp, BP);
        builds.push_back(build);
    }
    std::vector<ChunkPtr> probes;
    std::multiset<std::tuple<int32_t, int64_t, int64_t>> want;
    for (int base = 0; base < 20000; base += 4096) {
        auto pk = std::make_shared<Int32Column>(SRB_TYPE_INT);
        auto px = std::make_shared<Int64Column>(SRB_TYPE_BIGINT);
        for (int i = base; i < std::min(base + 4096, 20000); i++) {
            const int32_t k = (int32_t)(splitmix(i + 9999999) % 4000);
            pk->append(k);
            px->append(i);
            auto range = table.equal_range(k);
            for (auto it = range.first; it != range.second; ++it) want.emplace(k, (int64_t)i, it->second);
        }
        auto c = std::make_shared<Chunk>();
        c->append_column(pk, PK);
        c->append_column(px, PX);
        probes.push_back(c);
    }
    JoinPlan jp;
    jp.build_key_slots = {BK};
    jp.probe_key_slots = {PK};
    jp.build_out_slots = {BP};
    jp.probe_out_slots = {PK, PX};
    auto joiner = std::make_shared<GpuHashJoiner>(ctx, jp);
    RuntimeState state(4096);
    ChunkSourceOperator bsrc(builds);
    GpuHashJoinBuildOperator bop(joiner);
    GpuHashJoinProbeOperator pop(joiner);
    CHECK(!pop.need_input()); // the probe side is blocked until the build side finished (hash_joiner.h:60-65)
    auto bres = run_pipeline(&state, {&bsrc, &bop});
    CHECK(bres.ok());
    if (!bres.ok()) std::printf("  build: %s\n", bres.status().to_string().c_str());
    ChunkSourceOperator psrc(probes);
    auto res = run_pipeline(&state, {&psrc, &pop});
    CHECK(res.ok());
    if (!res.ok()) {
        std::printf("  %s\n", res.status().to_string().c_str());
        return;
    }
    std::multiset<std::tuple<int32_t, int64_t, int64_t>> got;
    for (auto& ch : res.value()) {
        CHECK(ch->num_rows() <= 4096);
        const int32_t* k = (const int32_t*)ch->get_column_by_slot_id(PK)->raw_data();
        const int64_t* x = (const int64_t*)ch->get_column_by_slot_id(PX)->raw_data();
        const int64_t* p = (const int64_t*)ch->get_column_by_slot_id(BP)->raw_data();
        for (size_t r = 0; r < ch->num_rows(); r++) got.emplace(k[r], x[r], p[r]);
    }
    CHECK(got == want);
    CHECK(want.size() > 10000);
    std::printf("PASS inner join (%zu output rows in %zu chunks)\n", want.size(), res.value().size());
}

// ---- runtime filter: the build operator publishes it, the probe-side scan applies it, the probe sees far fewer rows -------------------
static void test_runtime_filter_scan(srb_ctx* ctx) {
    const int BK = 20, PK = 30, PX = 31;
    std::set<int32_t> keys;
    auto bk = std::make_shared<Int32Column>(SRB_TYPE_INT);
    for (int i = 0; i < 3000; i++) {
        const int32_t k = (int32_t)(splitmix(i) % 1000000);
        if (keys.insert(k).second) bk->append(k);
    }
    auto build = std::make_shared<Chunk>();
    build->append_column(bk, BK);
    std::vector<ChunkPtr> probes;
    std::multiset<std::pair<int32_t, int64_t>> want;
    const int n_probe = 40000;
    for (int base = 0; base < n_probe; base += 4096) {
        auto pk = std::make_shared<Int32Column>(SRB_TYPE_INT);
        auto px = std::make_shared<Int64Column>(SRB_TYPE_BIGINT);
        for (int i = base; i < std::min(base + 4096, n_probe); i++) {
            const int32_t k = (int32_t)(splitmix(i % 2 ? i + 77777 : i / 2) % 1000000); // every other row probes a build key
            pk->append(k);
            px->append(i);
            if (keys.count(k)) want.emplace(k, (int64_t)i);
        }
        auto c = std::make_shared<Chunk>();
        c->append_column(pk, PK);
        c->append_column(px, PX);
        probes.push_back(c);
    }
    JoinPlan jp;
    jp.build_key_slots = {BK};
    jp.probe_key_slots = {PK};
    jp.probe_out_slots = {PK, PX};
    auto joiner = std::make_shared<GpuHashJoiner>(ctx, jp);
    joiner->runtime_filter_key = 0;
    ScanAggPlan sp; // a scan without predicates of its own: only the runtime filter narrows it
    sp.out_slots = {PK, PX};
    GpuScanAggregateOperator scan(ctx, sp);
    scan.wait_for_runtime_filter(joiner, PK);
    RuntimeState state(4096);
    CHECK(scan.prepare(&state).ok());
    CHECK(!scan.need_input()); // blocked until the build side has published its filter
    ChunkSourceOperator bsrc({build});
    GpuHashJoinBuildOperator bop(joiner);
    CHECK(run_pipeline(&state, {&bsrc, &bop}).ok());
    CHECK(joiner->runtime_filter() != nullptr);
    int64_t nb = 0, mn = 0, mx = 0, ins = 0;
    int32_t hn = 0;
    CHECK(srb_runtime_filter_info(joiner->runtime_filter(), &nb, &mn, &mx, &hn, &ins) == SRB_OK);
    CHECK(ins == (int64_t)keys.size() && mn == *keys.begin() && mx == *keys.rbegin() && hn == 0 && nb == 128); // init(<=4096 keys): 2^(12-5) buckets
    CHECK(scan.need_input());
    ChunkSourceOperator psrc(probes);
    GpuHashJoinProbeOperator pop(joiner);
    auto res = run_pipeline(&state, {&psrc, &scan, &pop});
    CHECK(res.ok());
    if (!res.ok()) {
        std::printf("  %s\n", res.status().to_string().c_str());
        return;
    }
    std::multiset<std::pair<int32_t, int64_t>> got;
    for (auto& ch : res.value()) {
        const int32_t* k = (const int32_t*)ch->get_column_by_slot_id(PK)->raw_data();
        const int64_t* x = (const int64_t*)ch->get_column_by_slot_id(PX)->raw_data();
        for (size_t r = 0; r < ch->num_rows(); r++) got.emplace(k[r], x[r]);
    }
    CHECK(got == want);
    const int64_t kept = scan.rows_selected();
    CHECK(kept >= (int64_t)want.size() && kept < (int64_t)want.size() + n_probe / 20); // every match kept, nearly everything else dropped
    std::printf("PASS runtime filter (%zu matches, scan kept %lld of %d rows)\n", want.size(), (long long)kept, n_probe);
}

// ---- the error convention: statuses, never exceptions ----------------------------------------------------------------------------------
static void test_error_convention(srb_ctx* ctx) {
    ScanAggPlan plan;
    plan.nodes = {slot_ref(1, SRB_TYPE_BIGINT)};
    plan.aggs = {srb_agg_fn{SRB_AGG_SUM, 0}};
    GpuScanAggregateOperator op(ctx, plan);
    RuntimeState state;
    CHECK(op.prepare(&state).ok());
    CHECK(op.need_input() && !op.has_output() && !op.is_finished());
    auto v = std::make_shared<Int64Column>(SRB_TYPE_BIGINT, std::vector<int64_t>{1, 2, 3});
    auto c = std::make_shared<Chunk>();
    c->append_column(v, 2); // wrong slot
    Status st = op.push_chunk(&state, c);
    CHECK(!st.ok() && st.code() == SRB_INVALID_ARGUMENT && !st.message().empty());
    auto good = std::make_shared<Chunk>();
    good->append_column(v, 1);
    CHECK(op.push_chunk(&state, good).ok());
    CHECK(op.set_finishing(&state).ok());
    CHECK(op.set_finishing(&state).ok()); // idempotent (operator.h:66)
    CHECK(!op.push_chunk(&state, good).ok());
    CHECK(op.has_output());
    auto r = op.pull_chunk(&state);
    CHECK(r.ok() && r.value()->num_rows() == 1);
    auto eof = op.pull_chunk(&state);
    CHECK(!eof.ok() && eof.status().is_end_of_file()); // tolerated by the driver (pipeline_driver.cpp:329)
    CHECK(op.is_finished());
    RuntimeState cancelled;
    cancelled.cancel();
    GpuScanAggregateOperator op2(ctx, plan);
    CHECK(op2.prepare(&cancelled).ok());
    CHECK(op2.push_chunk(&cancelled, good).is_cancelled());
    std::printf("PASS error convention\n");
}

int main(int argc, char** argv) {
    srb_config cfg{};
    cfg.device = 0;
    cfg.chunk_size = 4096;
    cfg.world_size = 1;
    srb_ctx* ctx = nullptr;
    if (srb_ctx_create(&cfg, &ctx) != SRB_OK) {
        std::printf("no usable GPU: %s\n", srb_last_error());
        return argc > 1 && std::string(argv[1]) == "--allow-no-gpu" ? 0 : 2;
    }
    test_filter_and_decimal_sum(ctx);
    test_group_by_nullable_key(ctx);
    test_inner_join(ctx);
    test_runtime_filter_scan(ctx);
    test_error_convention(ctx);
    srb_ctx_destroy(ctx);
    std::printf("%d checks, %d failed\n", g_checks, g_failed);
    return g_failed ? 1 : 0;
}
