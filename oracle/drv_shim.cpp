// drv_shim.cpp — TEST INFRASTRUCTURE ONLY (LD_PRELOAD for the reference's EAGLE command-line drivers, built by oracle/Makefile into
// oracle/_ref/).  In this fork llama_output_reserve creates the logits buffer only when cparams.embeddings is set
// (R/src/llama-context.cpp:898-902) and the drivers crash in their first decode without it, but `--embeddings` is not an option of
// LLAMA_EXAMPLE_SPECULATIVE (R/common/arg.cpp).  This shim interposes common_context_params_to_llama (R/common/common.cpp:1275), calls
// the real one and switches embeddings on with pooling NONE — what SURVEY F6 prescribes for any harness over this fork.  Nothing of the
// reference is copied or modified; the product never loads this file.
#include "common.h"
#include <dlfcn.h>

struct llama_context_params common_context_params_to_llama(const common_params & params) {
    using fn_t = llama_context_params (*)(const common_params &);
    static fn_t real = (fn_t) dlsym(RTLD_NEXT, "_Z30common_context_params_to_llamaRK13common_params");
    llama_context_params c = real(params);
    c.embeddings   = true;
    c.pooling_type = LLAMA_POOLING_TYPE_NONE;
    return c;
}
