// MOCK of the few Dr.Jit names the merl / customized_measurment plugin hosts touch.
// Not Dr.Jit: no tracing, no JIT, no AD.  A CUDAArray<T> here is an eagerly materialised
// device buffer (cudaMalloc) so that plugins/mitsuba3/*.cpp can be compiled AND run against
// libmerl_b200 in this repository, where neither Mitsuba 3 nor Dr.Jit exist
// (SURVEY.md §7 item 8, H3).  Names follow Dr.Jit 1.x.
#pragma once
#include <cuda_runtime_api.h>

#include <algorithm>
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
        mock_cuda_check(cudaMemcpy(a.raw(), host, n * sizeof(T), cudaMemcpyHostToDevice), "cudaMemcpy H2D");
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
        mock_cuda_check(cudaMemcpy(host, raw(), n * sizeof(T), cudaMemcpyDeviceToHost), "cudaMemcpy D2H");
    }
    T* data();                                                      // raises while "recording" (below)
    const T* data() const;
    size_t size() const { return n; }
    T* raw() { return static_cast<T*>(buf.get()); }                 // mock internals: no symbolic check
    const T* raw() const { return static_cast<const T*>(buf.get()); }
};

// MOCK of an llvm_* variant's array: a packet in HOST memory (upstream: LLVMArray<T>, evaluated by
// LLVM-compiled kernels on the CPU).  Same member names as the CUDA mock so that the helpers below
// serve both; data() is a host pointer, which is what the plugin hands to the *_host entry points.
template <typename T> struct LLVMArray {
    using Value = T;
    std::shared_ptr<T[]> buf;
    size_t n = 0;
    LLVMArray() = default;
    LLVMArray(T v) { *this = full_(v, 1); }
    static LLVMArray empty_(size_t n) {
        LLVMArray a;
        a.n = n;
        a.buf = std::shared_ptr<T[]>(new T[n ? n : 1]());
        return a;
    }
    static LLVMArray full_(T v, size_t n) {
        LLVMArray a = empty_(n);
        for (size_t i = 0; i < n; ++i) a.buf[i] = v;
        return a;
    }
    static LLVMArray load_(const void* host, size_t n) {
        LLVMArray a = empty_(n);
        std::memcpy(a.buf.get(), host, n * sizeof(T));
        return a;
    }
    void store_(void* host) const { std::memcpy(host, buf.get(), n * sizeof(T)); }
    T* data();
    const T* data() const;
    size_t size() const { return n; }
};

// ---- JIT flags (drjit-core: jit_flag / jit_set_flag).  Recording is set WHILE Dr.Jit records a
// virtual function call or a loop symbolically (the default mode of cuda_* / llvm_* variants):
// variables are placeholders then, and evaluating one or taking its data() raises.  The mock
// reproduces that so the plugin's guard against it can be tested.
}  // namespace drjit
enum class JitFlag : uint32_t { LoopRecord = 1, VCallRecord = 4, Recording = 128 };   // drjit-core: global scope
inline uint32_t& mock_jit_flags() { static uint32_t f = 0; return f; }
inline int jit_flag(JitFlag f) { return (mock_jit_flags() & (uint32_t) f) != 0; }
inline void jit_set_flag(JitFlag f, int on) {
    if (on) mock_jit_flags() |= (uint32_t) f; else mock_jit_flags() &= ~(uint32_t) f;
}
namespace drjit {
inline void mock_no_symbolic(const char* what) {
    if (jit_flag(JitFlag::Recording))
        throw std::runtime_error(std::string("drjit mock: ") + what + " of a symbolic (placeholder) variable while recording");
}
// Dr.Jit's LLVM backend runs dr::eval()'s kernels asynchronously on its thread pool: the host memory
// behind data() is valid only after dr::sync_thread().  The mock computes nothing asynchronously, so it
// models the hazard instead: eval() marks host packets as in flight, sync_thread() clears the mark, and
// reading an LLVM array's data() in between raises.
inline bool& mock_llvm_in_flight() { static bool f = false; return f; }
inline int& mock_sync_thread_calls() { static int n = 0; return n; }
inline void mock_llvm_synced(const char* what) {
    if (mock_llvm_in_flight())
        throw std::runtime_error(std::string("drjit mock: ") + what + " of an LLVM array after dr::eval() without dr::sync_thread(): "
                                 "the LLVM backend may still be writing the packet");
}
template <typename T> T* LLVMArray<T>::data() { mock_no_symbolic("data()"); mock_llvm_synced("data()"); return buf.get(); }
template <typename T> const T* LLVMArray<T>::data() const { mock_no_symbolic("data()"); mock_llvm_synced("data()"); return buf.get(); }

template <typename T> struct is_cuda : std::false_type {};
template <typename T> struct is_cuda<CUDAArray<T>> : std::true_type {};
template <typename T> constexpr bool is_cuda_v = is_cuda<T>::value;
template <typename T> struct is_llvm : std::false_type {};
template <typename T> struct is_llvm<LLVMArray<T>> : std::true_type {};
template <typename T> constexpr bool is_llvm_v = is_llvm<T>::value;
template <typename T> constexpr bool is_jit_v = is_cuda_v<T> || is_llvm_v<T>;
template <typename T> T* CUDAArray<T>::data() { mock_no_symbolic("data()"); return raw(); }
template <typename T> const T* CUDAArray<T>::data() const { mock_no_symbolic("data()"); return raw(); }
template <typename T> constexpr bool is_diff_v = false;            // no AD in the mock

template <typename T, typename U> struct replace_scalar { using type = U; };
template <typename T, typename U> struct replace_scalar<CUDAArray<T>, U> { using type = CUDAArray<U>; };
template <typename T, typename U> struct replace_scalar<LLVMArray<T>, U> { using type = LLVMArray<U>; };
template <typename T> using mask_t = typename replace_scalar<T, bool>::type;
template <typename T> using uint32_array_t = typename replace_scalar<T, uint32_t>::type;

template <typename A> A empty(size_t n = 1) {
    if constexpr (is_jit_v<A>) return A::empty_(n); else return A{};
}
template <typename A> A zeros(size_t n = 1) {
    if constexpr (is_cuda_v<A>) {
        A a = A::empty_(n);
        mock_cuda_check(cudaMemset(a.raw(), 0, n * sizeof(typename A::Value)), "cudaMemset");
        return a;
    } else if constexpr (is_llvm_v<A>) return A::empty_(n);
    else return A(0);
}
template <typename A, typename V> A full(V v, size_t n = 1) {
    if constexpr (is_jit_v<A>) return A::full_((typename A::Value) v, n); else return A(v);
}
template <typename A> A load(const void* host, size_t n) { return A::load_(host, n); }
template <typename A> A map(void* device_ptr, size_t n) { return A::map_(device_ptr, n); }
template <typename... Ts> bool grad_enabled(const Ts&...) { return false; }   // no AD in the mock
template <typename A> const A& detach(const A& a) { return a; }
template <typename A> void store(void* host, const A& a) { a.store_(host); }

template <typename A> size_t width(const A& a) {
    if constexpr (is_jit_v<A>) return a.size(); else return 1;
}
template <typename A, typename... Ts> size_t width(const A& a, const Ts&... ts) {
    size_t w = width(a), r = width(ts...);
    return w > r ? w : r;
}
// Dr.Jit: dr::resize(a, n) broadcasts a width-1 array in place.
template <typename A> void resize(A& a, size_t n) {
    if constexpr (is_jit_v<A>) {
        if (a.size() == n) return;
        if (a.size() != 1) throw std::runtime_error("drjit mock: resize of a non-unit-width array");
        typename A::Value v;
        a.store_(&v);
        a = A::full_(v, n);
    }
}
// mock arrays are always materialised; while "recording" an evaluation raises, as in Dr.Jit
template <typename... Ts> void eval(const Ts&...) { mock_no_symbolic("eval()"); mock_llvm_in_flight() = true; }
// dr::select on host packets (the llvm branch of the plugin masks its results with it)
template <typename T> LLVMArray<T> select(const LLVMArray<bool>& m, const LLVMArray<T>& a, const LLVMArray<T>& b) {
    const size_t n = std::max(m.size(), std::max(a.size(), b.size()));
    LLVMArray<T> r = LLVMArray<T>::empty_(n);
    for (size_t i = 0; i < n; ++i)
        r.buf[i] = m.buf[m.size() == 1 ? 0 : i] ? a.buf[a.size() == 1 ? 0 : i] : b.buf[b.size() == 1 ? 0 : i];
    return r;
}
inline void sync_thread() {
    mock_cuda_check(cudaDeviceSynchronize(), "cudaDeviceSynchronize");
    mock_llvm_in_flight() = false;
    ++mock_sync_thread_calls();
}

}  // namespace drjit

// drjit-core C API: the CUDA stream of the calling thread's JIT state.
inline void* jit_cuda_stream() { return nullptr; }                 // mock: the default stream
