// ulvt/finite_fields/risc0_baby_bear.h -- the reference keeps risc0::Fp in this header (risc0_baby_bear.h:40-187) and aliases
// it in baby_bear.cuh; here the class lives in baby_bear.cuh and this name forwards to it.
#pragma once
#include "baby_bear.cuh"
