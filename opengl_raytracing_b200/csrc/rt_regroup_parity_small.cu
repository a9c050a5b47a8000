// rt_regroup_parity_small.cu — the regrouping megakernel, parity math, small-scene kernel, 12 warps per CTA (see rt_regroup_tu.inc).
#define RT_PARITY 1
#define RT_VARIANT_NS parity_rg_small
#define RT_RG_TU parity_small
#define RT_RG_ONLY_BIG 0
#ifndef RT_RG_WARPS_SMALL
#define RT_RG_WARPS_SMALL 12
#endif
#undef RT_PK_WARPS
#define RT_PK_WARPS RT_RG_WARPS_SMALL
#include "rt_regroup_tu.inc"
