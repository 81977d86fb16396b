// Host-build stub (test tooling): only default_rng_t is needed by the headers we include.
#pragma once
#include <pcg32/pcg32.h>
namespace tcnn { using default_rng_t = pcg32; }
