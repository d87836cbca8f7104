// weti.cu — calc_weti2 on the device (R/form_02b_calc_weti.R:3-110).  Placeholder until the ordered
// multiple-flow-direction accumulation lands; fails loudly (there is no CPU fallback).
#include "litap_internal.h"

int run_weti(litap_grid*, double) {
  return set_error(LITAP_EARG, "litap_weti: not implemented yet in this build");
}
