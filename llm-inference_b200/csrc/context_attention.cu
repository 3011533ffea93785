// Prefill (context) attention without materialised scores; placeholder for now.
#include "common.cuh"
extern "C" int b200_context_attention(const void*, const void*, const void*, void*, const int*,
                                      const int*, const int*, int, int, int, int, int, int, int,
                                      int, void*) {
    return B200_EUNSUPPORTED;
}
