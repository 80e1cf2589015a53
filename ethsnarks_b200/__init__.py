"""ethsnarks_b200 — B200 (sm_100a) Groth16 proving backend behind ethsnarks' prover entry points.

csrc/      hand-written CUDA kernels + the C ABI (libb200snark.so, header in include/b200snark.h)
lib.py     ctypes binding of the C ABI
host.py    host-side mirror of the reference interface (prove, load_proving_key, get_domain, ProverContext, Config)
"""
from .lib import (Context, Bases, ProvingKey, B200Error, SO_PATH, SYMBOLS, FMT_MCL_NOZK, FMT_MCL_FULL, FMT_ALT_NOZK,
                  FMT_ALT_FULL, FMT_CACHE)
from .host import (Config, ProverContext, Protoboard, R1CS, prove, load_proving_key, get_domain, proof_to_json,
                   generate_r1cs_example_with_field_input, generate_r1cs_example_with_binary_input)
