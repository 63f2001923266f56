// Host-side glue shared by the drop-in shims (sketch_gpu.cpp, dist_gpu.cpp).
//
// The shims keep the reference's five hot signatures (src/sketch.h:62-66, src/dist.h:35-37) and are compiled
// AGAINST THE REFERENCE'S OWN HEADERS (-I<RabbitKSSD>/src): a maintainer drops the two .cpp files next to
// sketch.cpp / dist.cpp and links libkssd_b200.so.  Everything device-side goes through the
// C-ABI in include/kssd_b200.h; there is no CPU implementation of the hot loops in here.
#pragma once
#include <cstdio>
#include <cstdlib>
#include <string>

#include "kssd_b200.h"

namespace kssd_host {

// One context per process, on the device named by KSSD_B200_DEVICE (default 0).  The reference reports errors
// with a message on stderr and exit(1) (e.g. src/dist.cpp:41-43); the shims do the same with the library's text.
inline kssd_ctx *context() {
    static kssd_ctx *ctx = nullptr;
    if (!ctx) {
        const char *dev = getenv("KSSD_B200_DEVICE");
        int rc = kssd_b200_init(&ctx, dev ? atoi(dev) : 0);
        if (rc != KSSD_OK) {
            fprintf(stderr, "ERROR: libkssd_b200: %s\n", kssd_b200_last_error());
            exit(1);
        }
    }
    return ctx;
}

inline void check(int rc, const char *what) {
    if (rc != KSSD_OK) {
        fprintf(stderr, "ERROR: %s: %s\n", what, kssd_b200_last_error());
        exit(1);
    }
}

}  // namespace kssd_host
