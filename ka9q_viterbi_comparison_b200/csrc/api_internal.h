// api_internal.h -- pieces of api.cu that the other host translation units of the library share: the calling thread's
// error message behind b200v_last_error() and the policy constants of a codec.
#pragma once
#include "../../include/b200_viterbi.h"
#include "codec.h"

namespace b200v {
int api_fail(int code, const char* msg);     // records msg for b200v_last_error(), returns code
int api_fill_params(CodecParams& c, int codec, int K, int R, const int* poly, const b200v_wy_config* wy);
}  // namespace b200v
