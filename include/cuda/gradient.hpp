// Forwarding header: keeps the reference's include path "cuda/gradient.hpp" working.
// The declarations live in ../vip_b200.hpp (one header for the whole drop-in API).
#pragma once
#include "../vip_b200.hpp"
