/*
 * cstone_b200/device_vector.hpp — cstone::DeviceVector<T> with the reference's interface
 * [ref: cuda/device_vector.h:24-62]: device-resident, resize keeps the old content and grows geometrically, upload
 * from / download to std::vector.  Header-only over the C ABI's device-memory helpers; no CUDA headers needed.
 */
#pragma once

#include <cstddef>
#include <utility>
#include <type_traits>
#include <vector>

#include "box.hpp"

namespace cstone
{

template<class T>
class DeviceVector
{
public:
    using value_type = T;

    DeviceVector() = default;
    explicit DeviceVector(std::size_t n) { resize(n); }
    DeviceVector(std::size_t n, T init)
    {
        std::vector<T> h(n, init);
        *this = h;
    }
    DeviceVector(const DeviceVector& o)
    {
        resize(o.size_);
        detail::check(cstone_b200_device_memcpy(p_, o.p_, size_ * sizeof(T), 2), "DeviceVector copy");
    }
    DeviceVector(DeviceVector&& o) noexcept { swap(o); }
    DeviceVector(const std::vector<T>& h) { *this = h; }
    DeviceVector(const T* first, const T* last)
    {
        resize(std::size_t(last - first));
        detail::check(cstone_b200_device_memcpy(p_, first, size_ * sizeof(T), 0), "DeviceVector upload");
    }
    ~DeviceVector() { cstone_b200_device_free(p_); }

    T* data() { return p_; }
    const T* data() const { return p_; }
    std::size_t size() const { return size_; }
    bool empty() const { return size_ == 0; }
    std::size_t capacity() const { return cap_; }

    void reserve(std::size_t n)
    {
        if (n <= cap_) return;
        T* q = nullptr;
        detail::check(cstone_b200_device_malloc(reinterpret_cast<void**>(&q), n * sizeof(T)), "DeviceVector reserve");
        detail::check(cstone_b200_device_memcpy(q, p_, size_ * sizeof(T), 2), "DeviceVector reserve");
        cstone_b200_device_free(p_);
        p_   = q;
        cap_ = n;
    }

    //! content of [0, min(size, n)) is kept; new elements are uninitialised, as in the reference
    void resize(std::size_t n)
    {
        if (n > cap_) reserve(std::size_t(double(n) * 1.01) + 1);
        size_ = n;
    }

    DeviceVector& swap(DeviceVector& o) noexcept
    {
        std::swap(p_, o.p_);
        std::swap(size_, o.size_);
        std::swap(cap_, o.cap_);
        return *this;
    }
    friend void swap(DeviceVector& a, DeviceVector& b) noexcept { a.swap(b); }

    DeviceVector& operator=(const std::vector<T>& h)
    {
        resize(h.size());
        detail::check(cstone_b200_device_memcpy(p_, h.data(), size_ * sizeof(T), 0), "DeviceVector upload");
        return *this;
    }
    DeviceVector& operator=(DeviceVector o) noexcept { return swap(o); }

    //! download [ref: cuda/cuda_utils.hpp toHost]
    std::vector<T> toHost() const
    {
        std::vector<T> h(size_);
        detail::check(cstone_b200_device_memcpy(h.data(), p_, size_ * sizeof(T), 1), "DeviceVector download");
        return h;
    }

private:
    T* p_             = nullptr;
    std::size_t size_ = 0, cap_ = 0;
};

//! @brief pointer + length of a device array owned by a tree or a Domain, valid until its next update / sync
template<class T>
struct DeviceSpan
{
    const T* ptr;
    std::size_t n;
    const T* data() const { return ptr; }
    std::size_t size() const { return n; }
    //! copy to the host
    std::vector<std::remove_const_t<T>> toHost() const
    {
        std::vector<std::remove_const_t<T>> h(n);
        if (n) detail::check(cstone_b200_device_memcpy(h.data(), ptr, n * sizeof(T), 1), "DeviceSpan download");
        return h;
    }
};

template<class T>
std::vector<T> toHost(const DeviceVector<T>& v)
{
    return v.toHost();
}

} // namespace cstone

template<class T>
T* rawPtr(cstone::DeviceVector<T>& v)
{
    return v.data();
}
template<class T>
const T* rawPtr(const cstone::DeviceVector<T>& v)
{
    return v.data();
}
