// MOCK of the few Dr.Jit names the merl / customized_measurment plugin hosts touch.
// Not Dr.Jit: no tracing, no JIT, no AD.  A CUDAArray<T> here is an eagerly materialised
// device buffer (cudaMalloc) so that plugins/mitsuba3/*.cpp can be compiled AND run against
// libmerl_b200 in this repository, where neither Mitsuba 3 nor Dr.Jit exist
// (SURVEY.md §7 item 8, H3).  Names follow Dr.Jit 1.x.
#pragma once
#include <cuda_runtime_api.h>

#include <cstdint>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <vector>

namespace drjit {

inline void mock_cuda_check(cudaError_t e, const char* what) {
    if (e != cudaSuccess) throw std::runtime_error(std::string("drjit mock: ") + what + ": " + cudaGetErrorString(e));
}

template <typename T> struct CUDAArray {
    using Value = T;
    std::shared_ptr<void> buf;
    size_t n = 0;

    CUDAArray() = default;
    CUDAArray(T v) { *this = full_(v, 1); }                       // literal: width-1 array
    static CUDAArray empty_(size_t n) {
        CUDAArray a;
        a.n = n;
        void* p = nullptr;
        mock_cuda_check(cudaMalloc(&p, (n ? n : 1) * sizeof(T)), "cudaMalloc");
        a.buf = std::shared_ptr<void>(p, [](void* q) { cudaFree(q); });
        return a;
    }
    static CUDAArray full_(T v, size_t n) {
        std::unique_ptr<T[]> h(new T[n ? n : 1]);                   // (std::vector<bool> has no data())
        for (size_t i = 0; i < n; ++i) h[i] = v;
        return load_(h.get(), n);
    }
    static CUDAArray load_(const void* host, size_t n) {
        CUDAArray a = empty_(n);
        mock_cuda_check(cudaMemcpy(a.data(), host, n * sizeof(T), cudaMemcpyHostToDevice), "cudaMemcpy H2D");
        return a;
    }
    // dr::map<A>(ptr, n): an array over memory owned by someone else (upstream jit_var_mem_map)
    static CUDAArray map_(void* device_ptr, size_t n) {
        CUDAArray a;
        a.n = n;
        a.buf = std::shared_ptr<void>(device_ptr, [](void*) {});
        return a;
    }
    void store_(void* host) const {
        mock_cuda_check(cudaMemcpy(host, data(), n * sizeof(T), cudaMemcpyDeviceToHost), "cudaMemcpy D2H");
    }
    T* data() { return static_cast<T*>(buf.get()); }
    const T* data() const { return static_cast<const T*>(buf.get()); }
    size_t size() const { return n; }
};

template <typename T> struct is_cuda : std::false_type {};
template <typename T> struct is_cuda<CUDAArray<T>> : std::true_type {};
template <typename T> constexpr bool is_cuda_v = is_cuda<T>::value;
template <typename T> constexpr bool is_llvm_v = false;            // no LLVM backend in the mock
template <typename T> constexpr bool is_jit_v = is_cuda_v<T>;
template <typename T> constexpr bool is_diff_v = false;            // no AD in the mock

template <typename T, typename U> struct replace_scalar { using type = U; };
template <typename T, typename U> struct replace_scalar<CUDAArray<T>, U> { using type = CUDAArray<U>; };
template <typename T> using mask_t = typename replace_scalar<T, bool>::type;
template <typename T> using uint32_array_t = typename replace_scalar<T, uint32_t>::type;

template <typename A> A empty(size_t n = 1) {
    if constexpr (is_cuda_v<A>) return A::empty_(n); else return A{};
}
template <typename A> A zeros(size_t n = 1) {
    if constexpr (is_cuda_v<A>) {
        A a = A::empty_(n);
        mock_cuda_check(cudaMemset(a.data(), 0, n * sizeof(typename A::Value)), "cudaMemset");
        return a;
    } else return A(0);
}
template <typename A, typename V> A full(V v, size_t n = 1) {
    if constexpr (is_cuda_v<A>) return A::full_((typename A::Value) v, n); else return A(v);
}
template <typename A> A load(const void* host, size_t n) { return A::load_(host, n); }
template <typename A> A map(void* device_ptr, size_t n) { return A::map_(device_ptr, n); }
template <typename... Ts> bool grad_enabled(const Ts&...) { return false; }   // no AD in the mock
template <typename A> const A& detach(const A& a) { return a; }
template <typename A> void store(void* host, const A& a) { a.store_(host); }

template <typename A> size_t width(const A& a) {
    if constexpr (is_cuda_v<A>) return a.size(); else return 1;
}
template <typename A, typename... Ts> size_t width(const A& a, const Ts&... ts) {
    size_t w = width(a), r = width(ts...);
    return w > r ? w : r;
}
// Dr.Jit: dr::resize(a, n) broadcasts a width-1 array in place.
template <typename A> void resize(A& a, size_t n) {
    if constexpr (is_cuda_v<A>) {
        if (a.size() == n) return;
        if (a.size() != 1) throw std::runtime_error("drjit mock: resize of a non-unit-width array");
        typename A::Value v;
        a.store_(&v);
        a = A::full_(v, n);
    }
}
template <typename... Ts> void eval(const Ts&...) {}               // mock arrays are always materialised
inline void sync_thread() { mock_cuda_check(cudaDeviceSynchronize(), "cudaDeviceSynchronize"); }

}  // namespace drjit

// drjit-core C API: the CUDA stream of the calling thread's JIT state.
inline void* jit_cuda_stream() { return nullptr; }                 // mock: the default stream
