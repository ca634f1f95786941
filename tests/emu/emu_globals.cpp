// TEST INFRASTRUCTURE ONLY: the two globals api.cu owns in the real library
namespace b200amg { char g_err[512]; long long g_launches = 0; }
