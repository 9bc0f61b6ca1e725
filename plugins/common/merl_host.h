// merl_host.h -- what both plugin hosts share: an owning handle for a libmerl_b200 table and
// the translation of C-ABI status codes into the host renderer's error mechanism.
// No computation lives here or in the plugin sources: every query goes through
// include/merl_b200.h, and there is no CPU fallback (a missing device fails construction).
#pragma once
#include <merl_b200.h>

#include <string>
#include <utility>

namespace merl_host {

// RAII owner of a merl_table_t (the plugin's only state besides its Properties).
class Table {
public:
    Table() = default;
    Table(const Table&) = delete;
    Table& operator=(const Table&) = delete;
    Table(Table&& o) noexcept : m_h(std::exchange(o.m_h, nullptr)) {}
    Table& operator=(Table&& o) noexcept {
        if (this != &o) { reset(); m_h = std::exchange(o.m_h, nullptr); }
        return *this;
    }
    ~Table() { reset(); }
    void reset() { if (m_h) merl_table_destroy(m_h); m_h = nullptr; }
    merl_table_t** put() { reset(); return &m_h; }
    const merl_table_t* get() const { return m_h; }
    merl_table_t* get() { return m_h; }
    merl_table_info_t info() const {
        merl_table_info_t i{};
        merl_table_info(m_h, &i);
        return i;
    }
private:
    merl_table_t* m_h = nullptr;
};

// `fail` is the host's thrower: Throw(...) for Mitsuba 3, Log(EError, ...) for Mitsuba 0.6.
template <typename Fail> inline void check(int rc, const char* plugin, const char* what, Fail&& fail) {
    if (rc != MERL_OK) fail(std::string(plugin) + ": " + what + " failed: " + merl_last_error());
}

}  // namespace merl_host
