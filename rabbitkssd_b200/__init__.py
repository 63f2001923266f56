"""rabbitkssd_b200: B200 (sm_100a) implementation of RabbitKSSD's sketch and alldist/dist hot paths.

csrc/       hand-written CUDA kernels + the C-ABI (include/kssd_b200.h) -> libkssd_b200.so
api.py      host-side Python mirror of the path over the C-ABI (tests, bench)
synth.py    deterministic synthetic genomes / sketches
formats.py  the reference's frozen file formats (.shuf/.sketch/.dict/.index/.alldist)
There is no CPU fallback: importing api and creating a Context without the built library or a GPU raises.
"""
