// ctc_misc.cu -- greedy path and forced alignment: not built yet.
#include "common.cuh"
using namespace b200;
#define STUB(name) do { set_error(name " is not implemented yet"); return B200_ERR_UNSUPPORTED_SHAPE; } while (0)
extern "C" int b200_ctc_greedy(const void*, int, int64_t, int64_t, int64_t, int64_t, int64_t, int64_t*, void*) { STUB("b200_ctc_greedy"); }
extern "C" size_t b200_ctc_viterbi_work_bytes(int64_t, int64_t) { return 0; }
extern "C" int b200_ctc_viterbi(const float*, int64_t, int64_t, const int64_t*, int64_t, int, int64_t*, void*, size_t, void*) { STUB("b200_ctc_viterbi"); }
