"""me_b200 - Python harness around the B200 open-loop ME engine (ctypes over the C-ABI).

abi      ctypes mirror of include/me_gpu.h
planes   padded picture planes and result slabs
synth    synthetic sequences and the random-access prediction structure
engine   the product: thin wrapper over libme_gpu.so (CUDA); no CPU fallback
(the wrappers of the CPU oracle and of the compiled reference are test infrastructure and live under tests/)
"""
