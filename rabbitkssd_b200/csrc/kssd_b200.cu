// Unity translation unit of libkssd_b200.so: the kernels launch each other's helpers, and a single TU avoids
// relocatable device code.  Build: see rabbitkssd_b200/build.py (nvcc -gencode arch=compute_100a,code=sm_100a).
#include "sketch.cu"
#include "radix.cu"
#include "segsort.cu"
#include "index.cu"
#include "hash64.cu"
#include "fasta.cu"
#include "fastq.cu"
#include "setops.cu"
#include "dist.cu"
#include "capi.cu"
#include "noalloc.cu"
