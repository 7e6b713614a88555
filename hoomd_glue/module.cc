// module.cc — the force-path slice of /root/reference/src/module.cc:110-166 for the B200 build. The export functions keep the
// reference's names, so the reference's module.cc works UNCHANGED; this file exists for builds that want the force path
// alone (no velocity computes / flow fields): it declares and calls exactly the exporters of the
// classes BASELINE.json names. CPU exporters come from the reference's own export_*.cc.inc, GPU exporters from this directory
// (pair) or the reference's (bond / dpd / aniso: their host classes are untouched — only their kernel drivers are replaced).
#include <pybind11/pybind11.h>

namespace hoomd
    {
namespace azplugins
    {
namespace detail
    {
// bond
void export_PotentialBondDoubleWell(pybind11::module&);
void export_PotentialBondQuartic(pybind11::module&);
// pair
void export_AnisoPotentialPairTwoPatchMorse(pybind11::module&);
void export_PotentialPairColloid(pybind11::module&);
void export_PotentialPairExpandedYukawa(pybind11::module&);
void export_PotentialPairHertz(pybind11::module&);
void export_PotentialPairPerturbedLennardJones(pybind11::module&);
// dpd
void export_PotentialPairDPDThermoGeneralWeight(pybind11::module&);
#ifdef ENABLE_HIP
void export_PotentialBondDoubleWellGPU(pybind11::module&);
void export_PotentialBondQuarticGPU(pybind11::module&);
void export_AnisoPotentialPairTwoPatchMorseGPU(pybind11::module&);
void export_PotentialPairColloidGPU(pybind11::module&);
void export_PotentialPairExpandedYukawaGPU(pybind11::module&);
void export_PotentialPairHertzGPU(pybind11::module&);
void export_PotentialPairPerturbedLennardJonesGPU(pybind11::module&);
void export_PotentialPairDPDThermoGeneralWeightGPU(pybind11::module&);
#endif
    } // namespace detail
    } // namespace azplugins
    } // namespace hoomd

PYBIND11_MODULE(_azplugins, m)
    {
    using namespace hoomd::azplugins::detail;
    export_PotentialBondDoubleWell(m);
    export_PotentialBondQuartic(m);
    export_AnisoPotentialPairTwoPatchMorse(m);
    export_PotentialPairColloid(m);
    export_PotentialPairExpandedYukawa(m);
    export_PotentialPairHertz(m);
    export_PotentialPairPerturbedLennardJones(m);
    export_PotentialPairDPDThermoGeneralWeight(m);
#ifdef ENABLE_HIP
    export_PotentialBondDoubleWellGPU(m);
    export_PotentialBondQuarticGPU(m);
    export_AnisoPotentialPairTwoPatchMorseGPU(m);
    export_PotentialPairColloidGPU(m);
    export_PotentialPairExpandedYukawaGPU(m);
    export_PotentialPairHertzGPU(m);
    export_PotentialPairPerturbedLennardJonesGPU(m);
    export_PotentialPairDPDThermoGeneralWeightGPU(m);
#endif
    }
