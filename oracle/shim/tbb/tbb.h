/* Test-infrastructure shim (oracle/): a SERIAL stand-in for the few TBB constructs the
 * reference's BVH builder uses (src/utils/bvh.cpp:54,116,191,218-232,350-352):
 * tbb::task with continuation-style spawning, parallel_reduce, parallel_for,
 * blocked_range, empty_task. Tasks run depth-first on the calling thread, ranges are
 * visited in ascending grain-sized chunks, so the build is deterministic (parallel_for's
 * partition becomes a stable partition). TBB is not installed in this image.
 * Not part of the product. */
#pragma once
#include <cstddef>
#include <vector>
#include <new>
#include "mutex.h"

namespace tbb {

template <typename T> class blocked_range {
public:
    blocked_range(T b, T e, size_t grain = 1) : b_(b), e_(e), g_(grain ? grain : 1) {}
    T begin() const { return b_; }
    T end() const { return e_; }
    size_t grainsize() const { return g_; }
private:
    T b_, e_; size_t g_;
};

template <typename Range, typename Body>
void parallel_for(const Range &r, const Body &body) {
    typedef decltype(r.begin()) T;
    for (T b = r.begin(); b < r.end();) {
        T e = (T) ((size_t) (r.end() - b) > r.grainsize() ? b + (T) r.grainsize() : r.end());
        body(Range(b, e, r.grainsize()));
        b = e;
    }
}

template <typename Range, typename Value, typename Map, typename Reduce>
Value parallel_reduce(const Range &r, const Value &identity, const Map &map, const Reduce &reduce) {
    typedef decltype(r.begin()) T;
    Value acc = identity;
    bool first = true;
    for (T b = r.begin(); b < r.end();) {
        T e = (T) ((size_t) (r.end() - b) > r.grainsize() ? b + (T) r.grainsize() : r.end());
        Value part = map(Range(b, e, r.grainsize()), identity);
        acc = first ? part : reduce(acc, part);
        first = false;
        b = e;
    }
    return acc;
}

struct task_alloc_tag {};

class task {
public:
    virtual ~task() {}
    virtual task *execute() = 0;

    static task_alloc_tag allocate_root() { return task_alloc_tag(); }
    task_alloc_tag allocate_continuation() { return task_alloc_tag(); }
    task_alloc_tag allocate_child() { return task_alloc_tag(); }
    void set_ref_count(int) {}
    void recycle_as_child_of(task &) {}
    static void spawn(task &t) { pending().push_back(&t); }

    static void spawn_root_and_wait(task &root) {
        std::vector<task *> &stack = pending();
        stack.push_back(&root);
        while (!stack.empty()) {
            task *t = stack.back();
            stack.pop_back();
            while (t) {
                task *next = t->execute();
                if (next != t) delete t;   /* returning `this` = recycled, keep running it */
                t = next;
            }
        }
    }
private:
    static std::vector<task *> &pending() { static thread_local std::vector<task *> s; return s; }
};

class empty_task : public task {
public:
    task *execute() override { return nullptr; }
};

class task_scheduler_init {
public:
    explicit task_scheduler_init(int = 0) {}
    static int default_num_threads() { return 1; }
};

} // namespace tbb

inline void *operator new(std::size_t sz, tbb::task_alloc_tag) { return ::operator new(sz); }
inline void operator delete(void *p, tbb::task_alloc_tag) { ::operator delete(p); }
