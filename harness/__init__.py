"""Test / bench harness shared by tests/, bench.py and __graft_entry__.smoke(): synthetic scenes,
ctypes bindings of the CPU oracle (oracle/libmsgs_oracle.so) and of the in-place reference CUDA build
(oracle/_ref/libref_rasterizer.so).  Nothing in here is on the product path."""
