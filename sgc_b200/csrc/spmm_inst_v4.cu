// Instantiations of the hop kernel for 4-float gathers (see spmm.cuh).  One translation unit per
// vector width so the variants compile in parallel.
#include "spmm.cuh"

namespace sgc {

#define SGC_HOP_SHAPES X(8,4,1) X(16,4,1) X(32,4,1) X(32,4,2)

hop_launch_fn find_hop_kernel_v4(int g, int u, bool fused, int stdmode, bool peers, bool hot)
{
#define X(G_, V_, U_)                                                                  \
    if (g == G_ && u == U_) {                                                          \
        if (peers || hot) {                                                            \
            if (stdmode != 0) return nullptr;                                          \
            if (peers && hot)                                                          \
                return fused ? &launch_hop<G_, V_, U_, true, 0, true, true>            \
                             : &launch_hop<G_, V_, U_, false, 0, true, true>;          \
            if (peers)                                                                 \
                return fused ? &launch_hop<G_, V_, U_, true, 0, true, false>           \
                             : &launch_hop<G_, V_, U_, false, 0, true, false>;         \
            return fused ? &launch_hop<G_, V_, U_, true, 0, false, true>               \
                         : &launch_hop<G_, V_, U_, false, 0, false, true>;             \
        }                                                                              \
        if (!fused && stdmode == 0) return &launch_hop<G_, V_, U_, false, 0, false, false>;          \
        if (fused && stdmode == 0) return &launch_hop<G_, V_, U_, true, 0, false, false>;            \
        if (fused && stdmode == 1) return &launch_hop<G_, V_, U_, true, 1, false, false>;            \
        if (fused && stdmode == 2) return &launch_hop<G_, V_, U_, true, 2, false, false>;            \
        if (!fused && stdmode == 1) return &launch_hop<G_, V_, U_, false, 1, false, false>;          \
        if (!fused && stdmode == 2) return &launch_hop<G_, V_, U_, false, 2, false, false>;          \
        return nullptr;                                                                \
    }
    SGC_HOP_SHAPES
#undef X
    return nullptr;
}

}  // namespace sgc
