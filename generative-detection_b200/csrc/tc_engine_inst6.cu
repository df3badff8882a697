#include "tc_engine_inst.cuh"
ODVAE_TC_ENGINE_INSTANCE(6)
