// Test-infrastructure stub (NOT product code): a pass-through stand-in for
// OpenImageDenoise so the reference's OpenPGL sources compile on a box that has
// no ISPC/OIDN.  "Filters" are identity: execute() leaves the in-place buffers
// untouched.  Only the ISGB *training* path would ever call execute(); the
// load-from-file path the render pass uses never does.
#pragma once
#include <cstddef>
#include <cstring>
#include <memory>
#include <vector>
namespace oidn {
enum class Format { Undefined = 0, Float = 1, Float2, Float3, Float4 };
struct BufferRef {
    std::shared_ptr<std::vector<char>> d;
    void write(size_t off, size_t n, const void *src) { std::memcpy(d->data() + off, src, n); }
    void read(size_t off, size_t n, void *dst) const { std::memcpy(dst, d->data() + off, n); }
};
struct FilterRef {
    void setImage(const char *, const BufferRef &, Format, size_t, size_t) {}
    void set(const char *, bool) {}
    void commit() {}
    void execute() {}
};
struct DeviceRef {
    void commit() {}
    BufferRef newBuffer(size_t n) { return BufferRef{std::make_shared<std::vector<char>>(n)}; }
    FilterRef newFilter(const char *) { return FilterRef{}; }
};
inline DeviceRef newDevice() { return DeviceRef{}; }
}  // namespace oidn
