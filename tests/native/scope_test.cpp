// Entry points that use scratch are serialised by a recursive lock held for the lifetime of their ArenaScope (csrc/context.cu;
// SURVEY 8(b) "Threading").  Needs no GPU: three threads open nested scopes; at most one thread may be inside at a time.
#include "common.cuh"
#include <thread>
#include <atomic>
#include <cstdio>
static int inside = 0; static std::atomic<int> violations{0};
static void worker(int iters) {
    for (int i = 0; i < iters; ++i) {
        aq::ArenaScope outer;
        if (++inside != 1) violations++;
        { aq::ArenaScope inner; { aq::ArenaScope inner2; } }
        if (--inside != 0) violations++;
    }
}
int main() {
    std::thread a(worker, 200000), b(worker, 200000), c(worker, 200000);
    a.join(); b.join(); c.join();
    printf("violations %d\n", violations.load());
    return violations.load() != 0;
}
