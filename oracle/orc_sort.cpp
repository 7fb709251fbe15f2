// orc_sort.cpp — ORDER BY ... [LIMIT] over a chunk: the reference's ChunksSorterTopn / ChunksSorterFullSort results.
// TEST INFRASTRUCTURE ONLY (see oracle/README.md): never linked or loaded by anything under starrocks_b200/.
//
// Follows be/src/exec/sorting/sorting.h:86-101 (SortDesc: sort_order = +1 / -1; NULLs come first in the OUTPUT when
// is_null_first, whatever the direction), be/src/exec/sorting/compare_column.cpp (per-column three-way compare, columns in
// order), be/src/exec/chunks_sorter_topn.cpp (rows [offset, offset + limit) of the sorted order) and
// be/src/exec/chunks_sorter_full_sort.cpp. Strings compare as Slice::compare (memcmp, then length); rows that tie on every key
// keep their arrival order (ChunksSorterTest.stable_sort). Pinned by the known answers of be/test/exec/chunks_sorter_test.cpp
// :703-747 (topn_sort_with_limit), :799-966 (full sorts by 2 / 3 / 4 columns), :969-1060 (offset + limit) in
// tests/test_oracle_golden.py.
#include <algorithm>
#include <cstring>
#include <numeric>
#include <stdexcept>

#include "orc_agg.h"

namespace orc {

namespace {
struct SortKey {
    const srb_column* col;
    int order;       // +1 ascending, -1 descending
    bool null_first; // NULLs first in the output
};

int compare_values(const srb_column& c, int64_t a, int64_t b) {
    if (is_binary(c.type)) {
        const char* base = (const char*)c.data;
        const uint32_t la = c.offsets[a + 1] - c.offsets[a], lb = c.offsets[b + 1] - c.offsets[b];
        const int r = memcmp(base + c.offsets[a], base + c.offsets[b], std::min(la, lb));
        if (r != 0) return r < 0 ? -1 : 1;
        return la < lb ? -1 : (la > lb ? 1 : 0);
    }
    if (c.type == SRB_TYPE_DOUBLE) {
        const double x = ((const double*)c.data)[a], y = ((const double*)c.data)[b];
        return x < y ? -1 : (x > y ? 1 : 0);
    }
    const int128_t x = load_int(c, a), y = load_int(c, b);
    return x < y ? -1 : (x > y ? 1 : 0);
}

// negative: row a comes before row b
int compare_rows(const std::vector<SortKey>& keys, int64_t a, int64_t b) {
    for (const SortKey& k : keys) {
        const bool na = k.col->nulls && k.col->nulls[a], nb = k.col->nulls && k.col->nulls[b];
        if (na || nb) {
            if (na && nb) continue;
            return (na == k.null_first) ? -1 : 1;
        }
        const int r = compare_values(*k.col, a, b) * k.order;
        if (r != 0) return r;
    }
    return 0;
}
} // namespace

Result* run_sort(const srb_chunk* in, int n_keys, const int32_t* key_slots, const uint8_t* is_asc, const uint8_t* nulls_first, int64_t offset,
                 int64_t limit) {
    std::vector<SortKey> keys;
    for (int k = 0; k < n_keys; k++) {
        const srb_column* c = find_slot(*in, key_slots[k]);
        if (!c) throw std::runtime_error("sort key slot not in chunk");
        keys.push_back({c, is_asc[k] ? 1 : -1, nulls_first[k] != 0});
    }
    std::vector<int64_t> perm((size_t)in->n_rows);
    std::iota(perm.begin(), perm.end(), 0);
    std::stable_sort(perm.begin(), perm.end(), [&](int64_t a, int64_t b) { return compare_rows(keys, a, b) < 0; });
    const int64_t first = std::min<int64_t>(std::max<int64_t>(offset, 0), in->n_rows);
    const int64_t last = limit < 0 ? in->n_rows : std::min<int64_t>(in->n_rows, first + limit);
    auto* r = new Result();
    r->cols.resize(in->n_cols);
    for (int c = 0; c < in->n_cols; c++) {
        const srb_column& s = in->cols[c];
        OutCol& o = r->cols[c];
        r->slot_ids.push_back(in->slot_ids[c]);
        o.type = s.type;
        o.precision = s.precision;
        o.scale = s.scale;
        o.has_nulls = s.nulls != nullptr;
        for (int64_t i = first; i < last; i++) {
            const int64_t row = perm[(size_t)i];
            if (s.nulls) o.nulls.push_back(s.nulls[row]);
            if (is_binary(s.type))
                o.push_bytes((const char*)s.data + s.offsets[row], s.offsets[row + 1] - s.offsets[row]);
            else {
                const int w = type_width(s.type);
                const size_t off = o.data.size();
                o.data.resize(off + w);
                memcpy(o.data.data() + off, (const uint8_t*)s.data + (size_t)w * row, w);
            }
        }
    }
    r->rows_in = in->n_rows;
    r->rows_selected = last - first;
    r->n_rows = last - first;
    r->finish();
    return r;
}

} // namespace orc
