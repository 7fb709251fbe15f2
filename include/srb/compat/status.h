// include/srb/compat/status.h — minimal mirror of the reference's Status / StatusOr (be/src/common/status.h:95-155,
// be/src/common/statusor.h) with the same names and meanings, so that the operator binding in srb/gpu_operators.h reads
// like the class a BE maintainer would add. Codes are TStatusCode values (gensrc/thrift/StatusCode.thrift:53-83), the
// same integers the C ABI returns.
#pragma once
#include <string>
#include <utility>

#include "../srb_c_api.h"

namespace starrocks {

class Status {
public:
    Status() = default;
    Status(int code, std::string msg) : _code(code), _msg(std::move(msg)) {}
    static Status OK() { return Status(); }
    static Status EndOfFile(std::string m) { return Status(SRB_END_OF_FILE, std::move(m)); }
    static Status InternalError(std::string m) { return Status(SRB_INTERNAL_ERROR, std::move(m)); }
    static Status InvalidArgument(std::string m) { return Status(SRB_INVALID_ARGUMENT, std::move(m)); }
    static Status NotSupported(std::string m) { return Status(SRB_NOT_IMPLEMENTED, std::move(m)); }
    static Status Cancelled(std::string m) { return Status(SRB_CANCELLED, std::move(m)); }
    static Status MemoryLimitExceeded(std::string m) { return Status(SRB_MEM_LIMIT_EXCEEDED, std::move(m)); }

    bool ok() const { return _code == SRB_OK; }
    bool is_end_of_file() const { return _code == SRB_END_OF_FILE; }
    bool is_cancelled() const { return _code == SRB_CANCELLED; }
    bool is_not_supported() const { return _code == SRB_NOT_IMPLEMENTED; }
    bool is_mem_limit_exceeded() const { return _code == SRB_MEM_LIMIT_EXCEEDED; }
    int code() const { return _code; }
    const std::string& message() const { return _msg; }
    std::string to_string() const { return ok() ? "OK" : ("[" + std::to_string(_code) + "] " + _msg); }

private:
    int _code = SRB_OK;
    std::string _msg;
};

template <typename T>
class StatusOr {
public:
    StatusOr(Status s) : _status(std::move(s)) {}   // NOLINT: implicit like the reference
    StatusOr(T v) : _value(std::move(v)) {}         // NOLINT
    bool ok() const { return _status.ok(); }
    const Status& status() const { return _status; }
    T& value() { return _value; }
    const T& value() const { return _value; }
    T& operator*() { return _value; }
    T* operator->() { return &_value; }

private:
    Status _status;
    T _value{};
};

#define SRB_RETURN_IF_ERROR(stmt)          \
    do {                                   \
        ::starrocks::Status _s = (stmt);   \
        if (!_s.ok()) return _s;           \
    } while (0)

} // namespace starrocks
