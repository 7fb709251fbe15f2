// include/srb/compat/column.h — minimal mirror of the reference's column classes, with the accessors the GPU operator
// binding needs: Column (be/src/column/column.h:41-480), FixedLengthColumnBase<T>::get_data()
// (column/fixed_length_column_base.h:213), DecimalV3Column<T> precision/scale (column/decimalv3_column.h:27-59),
// NullableColumn data/null columns, 1 = NULL (column/nullable_column.h:23-31), BinaryColumn bytes + uint32 offsets
// (column/binary_column.h:34-37). Storage is std::vector like the reference's Buffer<T> (column/vectorized_fwd.h:47).
#pragma once
#include <cstdint>
#include <memory>
#include <string>
#include <vector>

#include "../srb_c_api.h"

namespace starrocks {

using int128_t = __int128;

class Column {
public:
    virtual ~Column() = default;
    virtual size_t size() const = 0;
    virtual bool is_nullable() const { return false; }
    virtual bool is_binary() const { return false; }
    virtual int logical_type() const = 0; // SRB_TYPE_* == the reference's LogicalType values
    virtual int precision() const { return 0; }
    virtual int scale() const { return 0; }
    virtual const void* raw_data() const = 0;
};
using ColumnPtr = std::shared_ptr<Column>;

template <typename T>
class FixedLengthColumn : public Column {
public:
    explicit FixedLengthColumn(int logical_type) : _type(logical_type) {}
    FixedLengthColumn(int logical_type, std::vector<T> v) : _type(logical_type), _data(std::move(v)) {}
    size_t size() const override { return _data.size(); }
    int logical_type() const override { return _type; }
    const void* raw_data() const override { return _data.data(); }
    std::vector<T>& get_data() { return _data; }
    const std::vector<T>& get_data() const { return _data; }
    void append(const T& v) { _data.push_back(v); }

protected:
    int _type;
    std::vector<T> _data;
};
using Int32Column = FixedLengthColumn<int32_t>;
using Int64Column = FixedLengthColumn<int64_t>;
using UInt8Column = FixedLengthColumn<uint8_t>;
using NullColumn = UInt8Column;

template <typename T>
class DecimalV3Column final : public FixedLengthColumn<T> {
public:
    DecimalV3Column(int precision, int scale)
            : FixedLengthColumn<T>(sizeof(T) == 4 ? SRB_TYPE_DECIMAL32 : (sizeof(T) == 8 ? SRB_TYPE_DECIMAL64 : SRB_TYPE_DECIMAL128)),
              _precision(precision),
              _scale(scale) {}
    int precision() const override { return _precision; }
    int scale() const override { return _scale; }

private:
    int _precision, _scale;
};

class BinaryColumn final : public Column {
public:
    explicit BinaryColumn(int logical_type = SRB_TYPE_VARCHAR) : _type(logical_type), _offsets(1, 0) {}
    size_t size() const override { return _offsets.size() - 1; }
    bool is_binary() const override { return true; }
    int logical_type() const override { return _type; }
    const void* raw_data() const override { return _bytes.data(); }
    std::vector<uint8_t>& get_bytes() { return _bytes; }
    const std::vector<uint8_t>& get_bytes() const { return _bytes; }
    std::vector<uint32_t>& get_offset() { return _offsets; }
    const std::vector<uint32_t>& get_offset() const { return _offsets; }
    void append(const std::string& s) {
        _bytes.insert(_bytes.end(), s.begin(), s.end());
        _offsets.push_back((uint32_t)_bytes.size());
    }
    std::string get(size_t i) const { return std::string(_bytes.begin() + _offsets[i], _bytes.begin() + _offsets[i + 1]); }

private:
    int _type;
    std::vector<uint8_t> _bytes;
    std::vector<uint32_t> _offsets;
};

class NullableColumn final : public Column {
public:
    NullableColumn(ColumnPtr data, std::shared_ptr<NullColumn> nulls) : _data(std::move(data)), _nulls(std::move(nulls)) {}
    size_t size() const override { return _data->size(); }
    bool is_nullable() const override { return true; }
    bool is_binary() const override { return _data->is_binary(); }
    int logical_type() const override { return _data->logical_type(); }
    int precision() const override { return _data->precision(); }
    int scale() const override { return _data->scale(); }
    const void* raw_data() const override { return _data->raw_data(); }
    const ColumnPtr& data_column() const { return _data; }
    const std::shared_ptr<NullColumn>& null_column() const { return _nulls; }
    bool has_null() const {
        for (uint8_t b : _nulls->get_data())
            if (b) return true;
        return false;
    }
    bool is_null(size_t i) const { return _nulls->get_data()[i] != 0; }

private:
    ColumnPtr _data;
    std::shared_ptr<NullColumn> _nulls;
};

} // namespace starrocks
