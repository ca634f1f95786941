// TEST INFRASTRUCTURE ONLY: backing store for the `extern __shared__` array of the prepared row-pattern kernel
// (pyamg_b200/csrc/rowpat_ilp2.cu) when it is compiled for the host; blocks run one after the other.
namespace b200amg { alignas(16) unsigned char r2_smem[1 << 20]; }
