// TEST INFRASTRUCTURE ONLY: backing store for the `extern __shared__` array of the prepared row-pattern kernel
// (profiles/r01b/prepared/rowpat_ilp2.cu.txt) when it is compiled for the host; blocks run one after the other.
namespace b200amg { alignas(16) unsigned char r2_smem[1 << 20]; }
