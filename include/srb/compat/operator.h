// include/srb/compat/operator.h — minimal mirror of starrocks::pipeline::Operator (be/src/exec/pipeline/operator.h:43-344):
// same method names, argument meaning and return conventions (file:line beside each), plus the slice of PipelineDriver::process
// (be/src/exec/pipeline/pipeline_driver.cpp:246-472) that moves chunks between two adjacent operators.
#pragma once
#include <string>
#include <vector>

#include "chunk.h"
#include "status.h"

namespace starrocks {

class RuntimeState {
public:
    explicit RuntimeState(int chunk_size = 4096) : _chunk_size(chunk_size) {}
    int chunk_size() const { return _chunk_size; } // config::vector_chunk_size, common/config.h:776
    bool is_cancelled() const { return _cancelled; }
    void cancel() { _cancelled = true; }

private:
    int _chunk_size;
    bool _cancelled = false;
};

namespace pipeline {

class Operator {
public:
    explicit Operator(std::string name) : _name(std::move(name)) {}
    virtual ~Operator() = default;
    virtual Status prepare(RuntimeState* state) { return Status::OK(); }                   // operator.h:55
    virtual Status set_finishing(RuntimeState* state) { return Status::OK(); }             // :66 (idempotent)
    virtual Status set_finished(RuntimeState* state) { return Status::OK(); }              // :78
    virtual Status set_cancelled(RuntimeState* state) { return Status::OK(); }             // :84
    virtual void close(RuntimeState* state) {}                                             // :92
    virtual bool has_output() const = 0;                                                   // :95
    virtual bool need_input() const = 0;                                                   // :101
    virtual bool is_finished() const = 0;                                                  // :105
    virtual StatusOr<ChunkPtr> pull_chunk(RuntimeState* state) = 0;                        // :118
    virtual Status push_chunk(RuntimeState* state, const ChunkPtr& chunk) = 0;             // :121
    const std::string& get_name() const { return _name; }

protected:
    std::string _name;
};
using OperatorPtr = std::shared_ptr<Operator>;

// Source of pre-built chunks (stands in for the scan operator of a pipeline in the tests).
class ChunkSourceOperator final : public Operator {
public:
    explicit ChunkSourceOperator(std::vector<ChunkPtr> chunks) : Operator("chunk_source"), _chunks(std::move(chunks)) {}
    bool has_output() const override { return _pos < _chunks.size(); }
    bool need_input() const override { return false; }
    bool is_finished() const override { return _pos >= _chunks.size(); }
    StatusOr<ChunkPtr> pull_chunk(RuntimeState*) override { return _chunks[_pos++]; }
    Status push_chunk(RuntimeState*, const ChunkPtr&) override { return Status::InternalError("source operator"); }

private:
    std::vector<ChunkPtr> _chunks;
    size_t _pos = 0;
};

// The driver loop between adjacent operators, as PipelineDriver::process does it (pipeline_driver.cpp:286-373): move a chunk
// while the upstream has output and the downstream needs input; propagate finishing when the upstream is finished; tolerate
// EndOfFile from pull_chunk (:329); drop empty chunks (:340-342); fail on a pulled chunk larger than chunk_size (:345-351).
// Returns the chunks the last operator produced.
inline StatusOr<std::vector<ChunkPtr>> run_pipeline(RuntimeState* state, std::vector<Operator*> ops) {
    std::vector<ChunkPtr> out;
    for (auto* op : ops) SRB_RETURN_IF_ERROR(op->prepare(state));
    std::vector<bool> finishing(ops.size(), false);
    for (int guard = 0; guard < (1 << 28); guard++) {
        bool moved = false;
        for (size_t i = 0; i + 1 < ops.size(); i++) {
            Operator *cur = ops[i], *next = ops[i + 1];
            if (cur->has_output() && next->need_input()) {
                auto maybe = cur->pull_chunk(state);
                if (!maybe.ok() && !maybe.status().is_end_of_file()) return maybe.status();
                if (maybe.ok() && maybe.value() && !maybe.value()->is_empty()) {
                    if ((int)maybe.value()->num_rows() > state->chunk_size())
                        return Status::InternalError(cur->get_name() + " returned a chunk larger than chunk_size");
                    SRB_RETURN_IF_ERROR(next->push_chunk(state, maybe.value()));
                }
                moved = true;
            }
            if (cur->is_finished() && !finishing[i + 1]) {
                SRB_RETURN_IF_ERROR(next->set_finishing(state));
                finishing[i + 1] = true;
                moved = true;
            }
        }
        Operator* last = ops.back();
        if (last->has_output()) {
            auto maybe = last->pull_chunk(state);
            if (!maybe.ok() && !maybe.status().is_end_of_file()) return maybe.status();
            if (maybe.ok() && maybe.value() && !maybe.value()->is_empty()) {
                if ((int)maybe.value()->num_rows() > state->chunk_size())
                    return Status::InternalError(last->get_name() + " returned a chunk larger than chunk_size");
                out.push_back(maybe.value());
            }
            moved = true;
        }
        if (last->is_finished()) break;
        if (!moved) return Status::InternalError("pipeline is stuck: no operator can make progress");
    }
    for (auto* op : ops) op->close(state);
    return out;
}

} // namespace pipeline
} // namespace starrocks
