"""aquery2_b200 -- B200-native group-by / windowed aggregation behind AQuery++'s post-processor API.

Layout:  csrc/  CUDA kernels + the C ABI (include/aqgpu.h) -> libaqgpu.so
         host/  C++ drop-in headers with the reference's names (ColRef, vector_type, TableInfo, ...)
         aqgpu.py  ctypes plumbing for tests / bench;  synth.py  deterministic synthetic inputs;
         dist.py  multi-GPU partial-aggregate exchange over torch.distributed.
"""
