// d4b_fast.cu -- placeholder until the uniform-degree kernel lands.
#include "d4b_internal.h"
void d4b_fast_setup(d4b_mesh *, const d4b_mesh_desc_t *, const std::vector<double> &,
                    const std::vector<int32_t> &, std::vector<char> &) {}
void d4b_launch_fast(d4b_mesh *, const double *, double *) {}
