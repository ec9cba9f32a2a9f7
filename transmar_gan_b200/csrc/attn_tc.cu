// attn_tc.cu — placeholder translation unit (tensor-core attention kernel under construction)
#include "common.cuh"
#include "tc.cuh"
namespace b200mar {
B200MAR_DEFINE_TC_ERROR_GETTER(tc_error_attn)
}
