// rt_regroup_parity_big.cu — the regrouping megakernel, parity math, big-scene kernel, 20 warps per CTA (see rt_regroup_tu.inc).
#define RT_PARITY 1
#define RT_VARIANT_NS parity_rg_big
#define RT_RG_TU parity_big
#define RT_RG_ONLY_BIG 1
#ifndef RT_RG_WARPS_BIG
#define RT_RG_WARPS_BIG 20
#endif
#undef RT_PK_WARPS
#define RT_PK_WARPS RT_RG_WARPS_BIG
#include "rt_regroup_tu.inc"
