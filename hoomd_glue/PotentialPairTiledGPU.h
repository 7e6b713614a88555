// PotentialPairTiledGPU.h — the one host-side addition of the B200 build. HOOMD's PotentialPairGPU<E> (and its DPD / aniso
// siblings) hand the kernel driver nothing but pair_args_t, so the tile plans of the shared-memory-staged kernels live in
// libazb200.so's cache (azb_tile_cache_get, called by the kernel-driver stubs). The cache cannot see that HOOMD rewrote the
// list arrays IN PLACE; this subclass tells it, from HOOMD's own bookkeeping: NeighborList::hasBeenUpdated(timestep) is true
// exactly on the steps the list was rebuilt. Instantiated by export_PotentialPairGPU.cc.inc in place of PotentialPairGPU<E>
// (/root/reference/src/export_PotentialPairGPU.cc.inc:32); the Python-visible class name is unchanged.
// Base may be md::PotentialPairGPU<E>, md::PotentialPairDPDThermoGPU<E> or md::AnisoPotentialPairGPU<E>.
#pragma once
#include "azb200.h"
#include "hoomd/md/PotentialPairGPU.h"

namespace hoomd
    {
namespace azplugins
    {
template<class Base> class TiledGPU : public Base
    {
    public:
    using Base::Base;

    protected:
    void computeForces(uint64_t timestep) override
        {
        if (this->m_nlist->hasBeenUpdated(timestep))
            {
            ArrayHandle<unsigned int> d_nlist(this->m_nlist->getNListArray(), access_location::device, access_mode::read);
            azb_tile_cache_invalidate(d_nlist.data); // every slot layout built for this list
            }
        Base::computeForces(timestep);
        }
    };

template<class evaluator> using PotentialPairTiledGPU = TiledGPU<md::PotentialPairGPU<evaluator>>;
    } // namespace azplugins
    } // namespace hoomd
