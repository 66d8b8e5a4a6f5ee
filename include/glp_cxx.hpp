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
#include <cstdlib>

namespace glp
{
    // device used by the wrapper classes: GLP_DEVICE in the environment, else 0
    inline int default_device()
    {
        const char *d = std::getenv("GLP_DEVICE");
        return d ? std::atoi(d) : 0;
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
            check(glp_init(default_device(), &t.ctx), "glp_init");
        return t.ctx;
    }
} // namespace glp
#endif
