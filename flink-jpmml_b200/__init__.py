"""flink-jpmml_b200 -- B200-native batched PMML scoring behind flink-jpmml's prediction API.

Layout
  csrc/        PMML compiler (host C++), CUDA kernels (sm_100a), engine + C ABI  ->  libfjpmml.so
  capi.py      ctypes binding of include/fjpmml.h (what tests and bench.py drive)
  api.py       Python mirror of the reference's user-facing types, for tests that read like the reference's specs
  pmmlgen.py   synthetic PMML emitters for the benchmark configs
"""
__version__ = "0.1"
