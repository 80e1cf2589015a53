// TEST INFRASTRUCTURE ONLY (oracle/): empty stand-in so that the reference's src/ethsnarks.hpp, which includes the mcl
// curve header unconditionally (ethsnarks.hpp:4), compiles in the CURVE_ALT_BN128 configuration without mcl/gmpxx.
#pragma once
