// TEST INFRASTRUCTURE (oracle/_ref build only) -- not part of the product.
// Minimal stand-in for msgpack-c 6.1.0 (reference MODULE.bazel:24), which is
// neither vendored under /root/reference nor installed here. The reference's
// Detector.hh pulls in sift_cuda/perf/*.cuh, which only needs these names to
// *compile*; the dump path (Detector::setDataGen) is never enabled by us, so
// every function here is a no-op.
#pragma once
#include <cstddef>
#include <stdexcept>
#include <string>
#include <vector>

#define MSGPACK_DEFAULT_API_NS v1
#define MSGPACK_API_VERSION_NAMESPACE(ns) inline namespace ns
#define MSGPACK_DEFINE_MAP(...)

namespace msgpack {
MSGPACK_API_VERSION_NAMESPACE(MSGPACK_DEFAULT_API_NS) {

struct type_error : public std::runtime_error {
    type_error() : std::runtime_error("msgpack stub") {}
};

namespace type {
enum object_type { NIL = 0, ARRAY = 6 };
}

struct object;
struct object_array {
    unsigned size = 0;
    object* ptr = nullptr;
};

struct object {
    type::object_type type = type::NIL;
    union union_type {
        object_array array;
        union_type() : array() {}
    } via;
    template <typename T> T as() const { return T(); }
    template <typename T> void convert(T&) const {}
};

struct object_handle {
    object obj;
    const object& get() const { return obj; }
};

struct sbuffer {
    const char* data() const { return nullptr; }
    std::size_t size() const { return 0; }
    void write(const char*, std::size_t) {}
};

template <typename Stream>
struct packer {
    packer& pack_array(unsigned) { return *this; }
    packer& pack_float(float) { return *this; }
    template <typename T> packer& pack(const T&) { return *this; }
};

template <typename Stream, typename T>
inline void pack(Stream&, const T&) {}

inline object_handle unpack(const char*, std::size_t) { return object_handle(); }

namespace adaptor {
template <typename T, typename Enabler = void> struct pack;
template <typename T, typename Enabler = void> struct convert;
}

}  // inline namespace
}  // namespace msgpack
