// MOCK (compile-only guard) of the slice of hoomd/md/PotentialPairGPU.h + NeighborList.h + GlobalArray.h that
// hoomd_glue/PotentialPairTiledGPU.h touches: the virtual computeForces(timestep), m_nlist->hasBeenUpdated(timestep),
// m_nlist->getNListArray() and an ArrayHandle on it.
#pragma once
#include <memory>
#include <pybind11/pybind11.h>

#include "hoomd/HOOMDMath.h"
namespace hoomd
    {
struct access_location
    {
    enum Enum
        {
        host,
        device
        };
    };
struct access_mode
    {
    enum Enum
        {
        read,
        readwrite,
        overwrite
        };
    };
template<class T> class GlobalArray
    {
    public:
    T* m_d = nullptr;
    };
template<class T> class ArrayHandle
    {
    public:
    ArrayHandle(const GlobalArray<T>& a, access_location::Enum, access_mode::Enum) : data(a.m_d) { }
    T* const data;
    };
class SystemDefinition
    {
    };
namespace md
    {
class NeighborList
    {
    public:
    bool hasBeenUpdated(uint64_t) const
        {
        return true;
        }
    const GlobalArray<unsigned int>& getNListArray() const
        {
        return m_nlist;
        }

    private:
    GlobalArray<unsigned int> m_nlist;
    };
template<class evaluator> class PotentialPairGPU
    {
    public:
    PotentialPairGPU(std::shared_ptr<SystemDefinition>, std::shared_ptr<NeighborList> nlist) : m_nlist(nlist) { }
    virtual ~PotentialPairGPU() { }

    protected:
    virtual void computeForces(uint64_t timestep) { (void)timestep; }
    std::shared_ptr<NeighborList> m_nlist;
    };
namespace detail
    {
template<class T> void export_PotentialPairGPU(pybind11::module& m, const std::string& name)
    {
    pybind11::class_<PotentialPairGPU<T>, std::shared_ptr<PotentialPairGPU<T>>>(m, name.c_str())
        .def(pybind11::init<std::shared_ptr<SystemDefinition>, std::shared_ptr<NeighborList>>());
    }
    } // namespace detail
    } // namespace md
    } // namespace hoomd
