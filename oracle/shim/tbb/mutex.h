/* Test-infrastructure shim (oracle/): minimal tbb::mutex over std::mutex, enough for
 * nori/block.h:128,182 and src/utils/block.cpp:130,168 of the reference. TBB is not
 * installed in this image. Not part of the product. */
#pragma once
#include <mutex>
namespace tbb {
class mutex {
public:
    void lock() { m_.lock(); }
    void unlock() { m_.unlock(); }
    class scoped_lock {
    public:
        explicit scoped_lock(mutex &m) : m_(m) { m_.lock(); }
        ~scoped_lock() { m_.unlock(); }
        scoped_lock(const scoped_lock &) = delete;
        scoped_lock &operator=(const scoped_lock &) = delete;
    private:
        mutex &m_;
    };
    mutex() = default;
    /* ImageBlock is copied by value in a few places of the reference; a copied mutex is a fresh one */
    mutex(const mutex &) {}
    mutex &operator=(const mutex &) { return *this; }
private:
    std::mutex m_;
};
}
