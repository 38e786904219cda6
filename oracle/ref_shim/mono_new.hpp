// ORACLE -- TEST INFRASTRUCTURE ONLY.  See mono_new.cpp.
#pragma once
namespace refshim {
void monotonic_new_enable(int on);
int monotonic_new_enabled();
// RAII: while alive (and enabled), this thread's operator new returns strictly increasing addresses
struct MonoScope {
    MonoScope();
    ~MonoScope();
    MonoScope(const MonoScope&) = delete;
    MonoScope& operator=(const MonoScope&) = delete;
private:
    bool active_;
};
}  // namespace refshim
