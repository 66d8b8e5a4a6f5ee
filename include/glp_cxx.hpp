// C++ glue shared by the two drop-in headers (ntt_goldilocks.hpp, merkleTreeGL.hpp): a lazily created
// per-thread glp_ctx and the translation of GLP_E_* codes into the exceptions the reference throws.
//
// Why per thread: the reference uses DISTINCT NTT_Goldilocks / MerkleTreeGL objects concurrently from OpenMP
// threads (src/starkpil/fri/friProve.cpp:44-107) and constructs them where no GPU work should happen
// (src/starkpil/starks.hpp:81-82, config.generateProof() == false).  A glp_ctx owns one stream and is not
// thread-safe, so every host thread gets its own, created on the first call that needs the GPU.
#ifndef GLP_CXX_HPP
#define GLP_CXX_HPP
#include "glprover.h"
#include <stdexcept>
#include <string>
#include <vector>
#include <cstdlib>

namespace glp
{
    // device used by the wrapper classes: GLP_DEVICE in the environment, else 0
    inline int default_device()
    {
        const char *d = std::getenv("GLP_DEVICE");
        return d ? std::atoi(d) : 0;
    }

    // devices used by extendPol / merkelize: GLP_DEVICES="0,1,2,3" (a power-of-two count; the work is split over them,
    // include/glprover.h glp_multi_*), else the single device above
    inline std::vector<int> device_list()
    {
        std::vector<int> d;
        if (const char *e = std::getenv("GLP_DEVICES"))
        {
            const char *p = e;
            while (*p)
            {
                char *end = nullptr;
                const long v = std::strtol(p, &end, 10);
                if (end == p)
                    break;
                d.push_back((int)v);
                p = (*end == ',') ? end + 1 : end;
            }
        }
        if (d.empty())
            d.push_back(default_device());
        return d;
    }

    [[noreturn]] inline void raise(int rc, const char *where)
    {
        const std::string msg = std::string(where) + ": " + glp_last_error();
        if (rc == GLP_E_DOMAIN)
            throw std::range_error("Domain size too big for the curve");     // ntt_goldilocks.hpp:94-97
        if (rc == GLP_E_ARG)
            throw std::invalid_argument(msg);                                // the reference asserts (ntt_goldilocks.cpp:28)
        throw std::runtime_error(msg);                                       // CUDA / OOM: no CPU fallback, fail loudly
    }
    inline void check(int rc, const char *where)
    {
        if (rc != GLP_OK)
            raise(rc, where);
    }

    struct ThreadCtx
    {
        glp_ctx *ctx = nullptr;
        ~ThreadCtx()
        {
            if (ctx)
                glp_destroy(ctx);
        }
    };

    // the calling thread's context (created on first use; destroyed when the thread exits)
    inline glp_ctx *thread_ctx()
    {
        static thread_local ThreadCtx t;
        if (!t.ctx)
            check(glp_init(device_list()[0], &t.ctx), "glp_init");
        return t.ctx;
    }

    // GLP_DEVICES names more than one device: the big host-to-host calls (extendPol, merkelize) go through a glp_multi.
    // One per calling thread, created on first use -- Starks::genProof makes those calls from its main thread only
    // (src/starkpil/starks.cpp:53-57), the per-thread objects of friProve.cpp keep the single-device context above.
    inline bool multi_enabled()
    {
        static const bool on = device_list().size() > 1;
        return on;
    }
    struct ThreadMulti
    {
        glp_multi *m = nullptr;
        ~ThreadMulti()
        {
            if (m)
                glp_multi_destroy(m);
        }
    };
    inline glp_multi *thread_multi()
    {
        static thread_local ThreadMulti t;
        if (!t.m)
        {
            const std::vector<int> d = device_list();
            check(glp_multi_init((int)d.size(), d.data(), &t.m), "glp_multi_init");
        }
        return t.m;
    }
} // namespace glp
#endif
