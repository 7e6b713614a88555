// MOCK (compile-only guard) of hoomd/BoxDim.h: the accessors hoomd_glue/azb_glue.h uses.
#pragma once
#include "HOOMDMath.h"
namespace hoomd
    {
class BoxDim
    {
    public:
    BoxDim() : m_L {1, 1, 1}, m_xy(0), m_xz(0), m_yz(0), m_periodic {1, 1, 1} { }
    Scalar3 getL() const
        {
        return m_L;
        }
    Scalar getTiltFactorXY() const
        {
        return m_xy;
        }
    Scalar getTiltFactorXZ() const
        {
        return m_xz;
        }
    Scalar getTiltFactorYZ() const
        {
        return m_yz;
        }
    uchar3 getPeriodic() const
        {
        return m_periodic;
        }

    private:
    Scalar3 m_L;
    Scalar m_xy, m_xz, m_yz;
    uchar3 m_periodic;
    };
    } // namespace hoomd
