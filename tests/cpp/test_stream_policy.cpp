// Replays the device sink's AUTO policy (starrocks_b200/csrc/srb_stream_policy.h, no CUDA) against a model table on the CPU:
//   test_stream_policy <units> <rows_per_unit> <distinct: 1 = every key new, 0 = keys from a small domain> <l2_bytes>
// prints one line: units per row mode (aggregate / pass / selective), rows passed through, groups, final state.
// Model: a table of 32-byte slots that doubles (at least) whenever it is more than half full, as ScanAggOp::check_batch grows it.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <unordered_set>

#include "../../starrocks_b200/csrc/srb_stream_policy.h"

int main(int argc, char** argv) {
    if (argc < 5) return 2;
    const int64_t units = atoll(argv[1]), per = atoll(argv[2]);
    const bool distinct = atoi(argv[3]) != 0;
    srb::StreamPolicy pol;
    pol.l2_bytes = (size_t)atoll(argv[4]);
    std::unordered_set<uint64_t> table;
    uint64_t cap = 1024, next_key = 0, lcg = 12345;
    int64_t seen = 0, passed = 0, by_mode[3] = {0, 0, 0};
    for (int64_t u = 0; u < units; u++) {
        const srb::StreamRowMode mode = pol.next(per, table.size(), cap, cap * 32, seen, passed);
        by_mode[mode]++;
        int64_t hits = 0;
        for (int64_t r = 0; r < per; r++) {
            lcg = lcg * 6364136223846793005ULL + 1442695040888963407ULL;
            const uint64_t key = distinct ? next_key++ : (lcg >> 33) % 2000;
            const bool present = table.count(key) != 0;
            if (mode == srb::STREAM_ROWS_AGGREGATE) {
                hits += present;
                table.insert(key);
            } else if (mode == srb::STREAM_ROWS_SELECTIVE && present) {
                hits++;
            } else {
                passed++;
            }
        }
        while (table.size() * 2 > cap) cap *= 2;
        seen += per;
        pol.observe(hits, per, cap * 32);
    }
    printf("%lld %lld %lld %lld %zu %d\n", (long long)by_mode[0], (long long)by_mode[1], (long long)by_mode[2], (long long)passed, table.size(), (int)pol.state);
    return 0;
}
