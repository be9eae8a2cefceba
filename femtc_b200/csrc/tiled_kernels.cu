// Placeholder: the tiled TMA kernels land here; until then everything runs on the generic kernels.
#include "common.cuh"
namespace femtc
{
template <typename T>
long long launch_mass_tiled(const MassArgs<T>&) { return 0; }
template <typename T>
long long launch_stiffness_tiled(const StiffnessArgs<T>&) { return 0; }
int tiled_kernel_info(int, int, int, int, long long*) { return FEMTC_OK; }
template long long launch_mass_tiled<double>(const MassArgs<double>&);
template long long launch_mass_tiled<float>(const MassArgs<float>&);
template long long launch_stiffness_tiled<double>(const StiffnessArgs<double>&);
template long long launch_stiffness_tiled<float>(const StiffnessArgs<float>&);
} // namespace femtc
