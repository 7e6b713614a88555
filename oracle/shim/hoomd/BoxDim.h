// TEST INFRASTRUCTURE (oracle). Minimal stand-in for HOOMD-blue's hoomd/BoxDim.h so that the reference's barrier evaluator
// headers (/root/reference/src/PlanarBarrierEvaluator.h:8, SphericalBarrierEvaluator.h:8) compile unmodified on the host.
// HOOMD-blue v5 is not vendored under /root/reference; this restates only what those headers and the HarmonicBarrier
// loop (HarmonicBarrier.h:148-178) use: the origin-centred triclinic box with makeCoordinates, getNearestPlaneDistance and
// the one-image wrap. Written from the documented HOOMD v5 semantics (recalled, not copied): parity of these three members
// with HOOMD itself is unpinned, the reference's tests (test_external.py) use an orthorhombic box with particles inside it.
#ifndef AZB_ORACLE_SHIM_BOXDIM_H_
#define AZB_ORACLE_SHIM_BOXDIM_H_

#include "HOOMDMath.h"

namespace hoomd
    {
struct int3
    {
    int x, y, z;
    };
inline int3 make_int3(int x, int y, int z)
    {
    int3 v = {x, y, z};
    return v;
    }

class BoxDim
    {
    public:
    BoxDim(Scalar Lx, Scalar Ly, Scalar Lz, Scalar xy, Scalar xz, Scalar yz, bool px, bool py, bool pz)
        : m_xy(xy), m_xz(xz), m_yz(yz), m_px(px), m_py(py), m_pz(pz)
        {
        m_L = make_scalar3(Lx, Ly, Lz);
        m_lo = make_scalar3(-Lx / Scalar(2), -Ly / Scalar(2), -Lz / Scalar(2));
        m_hi = make_scalar3(m_lo.x + Lx, m_lo.y + Ly, m_lo.z + Lz);
        }

    //! fractional coordinates -> Cartesian
    Scalar3 makeCoordinates(const Scalar3& f) const
        {
        Scalar3 v = make_scalar3(m_lo.x + f.x * m_L.x, m_lo.y + f.y * m_L.y, m_lo.z + f.z * m_L.z);
        v.x += m_xy * v.y + m_xz * v.z;
        v.y += m_yz * v.z;
        return v;
        }

    //! distance between opposite faces
    Scalar3 getNearestPlaneDistance() const
        {
        const Scalar term = m_xy * m_yz - m_xz;
        return make_scalar3(m_L.x / fast::sqrt(Scalar(1) + m_xy * m_xy + term * term),
                            m_L.y / fast::sqrt(Scalar(1) + m_yz * m_yz),
                            m_L.z);
        }

    //! bring a position back by one image per periodic direction
    void wrap(Scalar3& w, int3& img) const
        {
        if (m_pz)
            {
            if (w.z >= m_hi.z)
                {
                w.z -= m_L.z;
                w.y -= m_L.z * m_yz;
                w.x -= m_L.z * m_xz;
                img.z++;
                }
            else if (w.z < m_lo.z)
                {
                w.z += m_L.z;
                w.y += m_L.z * m_yz;
                w.x += m_L.z * m_xz;
                img.z--;
                }
            }
        if (m_py)
            {
            const Scalar tilt_y = m_yz * w.z;
            if (w.y >= m_hi.y + tilt_y)
                {
                w.y -= m_L.y;
                w.x -= m_L.y * m_xy;
                img.y++;
                }
            else if (w.y < m_lo.y + tilt_y)
                {
                w.y += m_L.y;
                w.x += m_L.y * m_xy;
                img.y--;
                }
            }
        if (m_px)
            {
            const Scalar tilt_x = (m_xz - m_xy * m_yz) * w.z + m_xy * w.y;
            if (w.x >= m_hi.x + tilt_x)
                {
                w.x -= m_L.x;
                img.x++;
                }
            else if (w.x < m_lo.x + tilt_x)
                {
                w.x += m_L.x;
                img.x--;
                }
            }
        }

    private:
    Scalar3 m_L, m_lo, m_hi;
    Scalar m_xy, m_xz, m_yz;
    bool m_px, m_py, m_pz;
    };
    } // namespace hoomd
#endif
