// acs_grp_punct.cu -- the group kernels of acs_grp.cu instantiated with the punctured symbol fetch (depuncturing
// fused into the symbol load): the same source compiled with GRP_PUNCT_TU = 1, exporting launch_acs_grp_punct only.
#define GRP_PUNCT_TU 1
#include "acs_grp.cu"
