// srb_prog.h — the lowered, device-readable form of an srb_scan_agg_plan ("program"), shared by host and device.
//
// The host lowers the plan (flattened Expr tree + predicates + aggregate list, include/srb/srb_c_api.h) once per
// operator: slot ids become indexes into a bound-column array, decimal typing decisions (fast-mul, scale
// adjustment factors, overflow -> NULL) are resolved into explicit node fields, and the aggregate states get their
// byte offsets inside a group-table slot. Kernels either interpret this program (runtime form) or are instantiated
// on a compile-time copy of the same structures (static shapes, srb_shapes.cuh) so the interpreter folds away.
#pragma once

#include "srb_dev.cuh"

namespace srb {

constexpr int MAX_COLS = 16;
constexpr int MAX_NODES = 24;
constexpr int MAX_PREDS = 16;
constexpr int MAX_KEYS = 8;
constexpr int MAX_AGGS = 12;
constexpr int MAX_KEY_WORDS = 8; // a packed group key is at most 64 bytes (one or two words: the fast paths; three to eight: "wide" keys)

// column kinds
constexpr int CK_INT = 0, CK_DOUBLE = 1, CK_BINARY = 2;

struct ColBind {
    const void* data;
    const uint8_t* nulls;    // may be null
    const uint32_t* offsets; // binary only
    int32_t width;           // 1,2,4,8,16 (0 for binary)
    int32_t kind;
};

// node ops (device level)
constexpr int DOP_LOAD = 1, DOP_LIT = 2, DOP_CAST = 3, DOP_ADD = 4, DOP_SUB = 5, DOP_MUL = 6, DOP_MULFAST = 7;

struct DNode {
    int32_t op;
    int32_t width;   // result width in bytes (1,2,4,8,16)
    int32_t lhs, rhs;
    int32_t col;     // DOP_LOAD
    int32_t checked; // decimal semantics: overflow -> NULL; otherwise wraps in `width`
    int32_t adj_l, adj_r; // ADD/SUB: operand must be multiplied (checked) by fl / fr first
    int32_t kind;    // CK_* of the value
    int32_t _pad;
    uint64_t lit_lo; // LIT value / CAST factor
    int64_t lit_hi;
    uint64_t fl_lo; // scale factor of the left operand
    int64_t fl_hi;
    uint64_t fr_lo;
    int64_t fr_hi;
};

struct DPred {
    int32_t col;
    int32_t op;    // SRB_PRED_*
    int32_t group; // 0 = conjunct; otherwise OR-group id (groups are contiguous after lowering)
    int32_t lit_len; // binary columns: bytes of the literal, which sit in (lit_lo, lit_hi) in memory order
    uint64_t lit_lo;
    int64_t lit_hi;
};

// aggregate functions, same codes as SRB_AGG_*
constexpr int FN_SUM = 1, FN_AVG = 2, FN_COUNT = 3, FN_MIN = 4, FN_MAX = 5;

struct DAgg {
    int32_t fn;
    int32_t arg;       // node index, -1 = count(*)
    int32_t is_double; // state holds a double in `lo`
    int32_t avg_as_double; // AVG over plain integers accumulates in double (avg.h:49-53)
    int32_t merge;     // 1: the argument is an intermediate state column (phase-2 merge)
    int32_t blob_col;  // merge of AVG: bound column index of the VARBINARY {sum,count} blobs
    int32_t off_lo, off_hi, off_n; // byte offsets inside a slot; -1 = field not kept
    int32_t wide;      // MIN/MAX compares 128-bit (lo,hi) pairs
};

struct DKey {
    int32_t node;     // must be a DOP_LOAD node
    int32_t nullable; // contributes a leading null byte
    int32_t width;    // fixed width; 0 = binary ([u32 len][bytes])
    int32_t _pad;
};

struct DProg {
    int32_t n_cols, n_nodes, n_preds, n_keys, n_aggs;
    int32_t phase_seed;  // 1 or 2: which phmap seed the table hash uses (agg_hash_variant.h:303-308)
    int32_t key_bytes;   // fixed packed key length, 0 when a binary key makes it per-row
    int32_t key_words;   // 1 or 2 (8- or 16-byte key area in a slot), 3..MAX_KEY_WORDS for wide keys (the words past the second follow pair B)
    int32_t slot_bytes;  // stride of one table slot (multiple of 16)
    int32_t off_state;   // first byte after the key header (16, 32, or 32 + the wide key's further words rounded up to 16)
    int32_t _pad[2];
    ColBind cols[MAX_COLS];
    DNode nodes[MAX_NODES];
    DPred preds[MAX_PREDS];
    DKey keys[MAX_KEYS];
    DAgg aggs[MAX_AGGS];
};

// one evaluated value in a row pipeline
struct Val {
    uint64_t lo;
    int64_t hi;
    bool null;
};

// Group-table slot: [A: meta u64 | key word 0] [B: key word 1 | ready flag  (16-byte keys only)] [states ...], 16-byte aligned.
// meta = (hash & ~3) | 3 once the slot is taken, 0 while empty. A is claimed with ONE 128-bit compare-and-swap that writes the
// tag and the first key word together, so an 8-byte key is published atomically: no lock state, no fence, no spinning. The
// second key word of a 16-byte key follows in pair B as a single 16-byte store whose flag makes it self-validating.
// Every aggregate state is valid when zero-filled (MIN/MAX keep order-preserving biased encodings), so a fresh slot needs
// no initialisation.
constexpr uint64_t SLOT_EMPTY = 0, SLOT_READY = 3;
constexpr int TABLE_MAX_PROBE = 256; // a longer linear probe means the table is too full: the row is deferred and replayed

struct TableView {
    uint8_t* slots;
    uint64_t mask; // capacity - 1
    int32_t slot_bytes;
    int32_t key_words;
};

} // namespace srb
