// include/srb/compat/chunk.h — minimal mirror of starrocks::Chunk (be/src/column/chunk.h:50-318): columns addressed by
// slot id through a slot_id -> index map (:314).
#pragma once
#include <map>
#include <memory>
#include <vector>

#include "column.h"

namespace starrocks {

using SlotId = int32_t;

class Chunk {
public:
    size_t num_rows() const { return _columns.empty() ? 0 : _columns[0]->size(); }
    size_t num_columns() const { return _columns.size(); }
    bool is_empty() const { return num_rows() == 0; }
    void append_column(ColumnPtr column, SlotId slot_id) {
        _slot_id_to_index[slot_id] = _columns.size();
        _slots.push_back(slot_id);
        _columns.push_back(std::move(column));
    }
    const ColumnPtr& get_column_by_slot_id(SlotId slot_id) const { return _columns[_slot_id_to_index.at(slot_id)]; }
    const ColumnPtr& get_column_by_index(size_t idx) const { return _columns[idx]; }
    bool is_slot_exist(SlotId slot_id) const { return _slot_id_to_index.count(slot_id) != 0; }
    const std::map<SlotId, size_t>& get_slot_id_to_index_map() const { return _slot_id_to_index; }
    const std::vector<SlotId>& slots_in_order() const { return _slots; }

private:
    std::vector<ColumnPtr> _columns;
    std::vector<SlotId> _slots;
    std::map<SlotId, size_t> _slot_id_to_index;
};
using ChunkPtr = std::shared_ptr<Chunk>;

} // namespace starrocks
