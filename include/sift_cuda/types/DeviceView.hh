// The reference exposes its results as thrust::device_vector members (Detector.hh:54-57). The engine owns
// one slab of device memory behind the C ABI, so the mirror exposes borrowed views with the part of the
// vector interface callers use (data(), size(), begin(), operator[] is deliberately absent: device memory).
#pragma once
#include <cstddef>
#include <type_traits>

namespace sift_cuda {

template <typename T>
class DeviceView {
  public:
    DeviceView() = default;
    DeviceView(const T* p, size_t n) : p_(p), n_(n) {}
    const T* data() const { return p_; }
    const T* begin() const { return p_; }
    const T* end() const { return p_ + n_; }
    size_t size() const { return n_; }
    bool empty() const { return n_ == 0; }

  private:
    const T* p_ = nullptr;
    size_t n_ = 0;
};

template <typename T>
inline const T* device_ptr(const DeviceView<T>& v) { return v.data(); }

// thrust::device_vector<T> (or anything else whose data() yields a thrust-style pointer with .get()).
template <class Vec>
inline auto device_ptr(const Vec& v) -> decltype(v.data().get()) { return v.data().get(); }

}  // namespace sift_cuda
