// ORACLE -- TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product path.
//
// A replacement for the global operator new / delete of the reference build (oracle/_ref/libgforb_ref*.so, linked
// -Bsymbolic so it binds inside that library only).  Why it exists: ORBextractor::DistributeOctTree sorts
// pair<int, ExtractorNode*> (src/ORBextractor.cc:591,684), so nodes holding the same number of keys are ordered by
// the HEAP ADDRESS of their std::list cell.  Under glibc malloc freed cells are recycled, the order depends on the
// allocator's history, and the reference returns different keypoints for the same image from one call to the next
// (tests/test_oracle_vs_ref.py::test_reference_is_allocator_dependent shows it).  Inside a "monotonic scope" this
// allocator hands out strictly increasing addresses and never recycles, which makes the tie rule a function of the
// input alone: "created later = higher address = split first".  That is the rule the restatement (and the CUDA path)
// froze; with it in force the reference's own code is reproducible and can be compared bit for bit.
// Outside a scope (the timed CPU arm, for one) everything goes to malloc / free as usual.
#include <sys/mman.h>

#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <new>

#include "mono_new.hpp"

namespace {

constexpr size_t kSlotBytes = (size_t)1 << 29;    // 512 MB of address space per scope (pages are touched on demand)
constexpr int kSlots = 32;

struct Slot {
    std::atomic<int> state{0};        // 0 free, 1 open (a scope is allocating), 2 closed with live objects
    std::atomic<long> live{0};
    size_t off = 0;
};

char* g_region = nullptr;
Slot g_slot[kSlots];
std::atomic<int> g_enabled{0};
thread_local Slot* t_slot = nullptr;
thread_local int t_depth = 0;

char* region() {
    static char* r = [] {
        void* p = mmap(nullptr, kSlotBytes * kSlots, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
        if (p == MAP_FAILED) { std::perror("mono_new: mmap"); std::abort(); }
        g_region = (char*)p;
        return (char*)p;
    }();
    return r;
}

void retire(Slot* s) {
    const int i = (int)(s - g_slot);
    madvise(g_region + kSlotBytes * (size_t)i, kSlotBytes, MADV_DONTNEED);
    s->off = 0;
    s->state.store(0, std::memory_order_release);
}

inline void* take(size_t n) {
    Slot* s = t_slot;
    if (s) {
        const size_t o = (s->off + 15) & ~(size_t)15;
        if (o + n > kSlotBytes) { std::fputs("mono_new: scope arena exhausted\n", stderr); std::abort(); }
        s->off = o + (n ? n : 1);
        s->live.fetch_add(1, std::memory_order_relaxed);
        return g_region + kSlotBytes * (size_t)(s - g_slot) + o;
    }
    void* p = std::malloc(n ? n : 1);
    if (!p) throw std::bad_alloc();
    return p;
}

inline void give(void* p) noexcept {
    if (!p) return;
    char* c = (char*)p;
    if (g_region && c >= g_region && c < g_region + kSlotBytes * kSlots) {
        Slot* s = &g_slot[(size_t)(c - g_region) / kSlotBytes];
        if (s->live.fetch_sub(1, std::memory_order_acq_rel) == 1 && s->state.load(std::memory_order_acquire) == 2) {
            int two = 2;
            if (s->state.compare_exchange_strong(two, 3)) retire(s);
        }
        return;
    }
    std::free(p);
}

}  // namespace

namespace refshim {

void monotonic_new_enable(int on) { g_enabled.store(on, std::memory_order_relaxed); }
int monotonic_new_enabled() { return g_enabled.load(std::memory_order_relaxed); }

MonoScope::MonoScope() : active_(false) {
    if (!g_enabled.load(std::memory_order_relaxed)) return;
    if (t_depth++ > 0) { active_ = true; return; }
    region();
    for (int i = 0; i < kSlots; ++i) {
        int zero = 0;
        if (g_slot[i].state.compare_exchange_strong(zero, 1)) {
            g_slot[i].off = 0;
            t_slot = &g_slot[i];
            active_ = true;
            return;
        }
    }
    std::fputs("mono_new: no free scope arena (objects of earlier scopes still alive)\n", stderr);
    std::abort();
}

MonoScope::~MonoScope() {
    if (!active_) return;
    if (--t_depth > 0) return;
    Slot* s = t_slot;
    t_slot = nullptr;
    s->state.store(2, std::memory_order_release);
    if (s->live.load(std::memory_order_acquire) == 0) {
        int two = 2;
        if (s->state.compare_exchange_strong(two, 3)) retire(s);
    }
}

}  // namespace refshim

void* operator new(size_t n) { return take(n); }
void* operator new[](size_t n) { return take(n); }
void* operator new(size_t n, const std::nothrow_t&) noexcept { try { return take(n); } catch (...) { return nullptr; } }
void* operator new[](size_t n, const std::nothrow_t&) noexcept { try { return take(n); } catch (...) { return nullptr; } }
void operator delete(void* p) noexcept { give(p); }
void operator delete[](void* p) noexcept { give(p); }
void operator delete(void* p, size_t) noexcept { give(p); }
void operator delete[](void* p, size_t) noexcept { give(p); }
void operator delete(void* p, const std::nothrow_t&) noexcept { give(p); }
void operator delete[](void* p, const std::nothrow_t&) noexcept { give(p); }
