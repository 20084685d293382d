// Host-side mechanism record produced by the mechanism compiler (mech_compiler.cpp).
// It plays the mechanism-load role of CanteraMixture in the reference
// (src/dfCanteraMixture/CanteraMixture.C:56-68: Cantera::newSolution(mechanismFile, ""))
// and holds the constants dfChemistryModel caches in its constructor
// (species order src/dfChemistryModel/dfChemistryModel.C:234-237, hc_ = Hf298SS/W :335-338).
//
// Everything is SI-kmol: m, kg, s, kmol, K, J  (Cantera's internal units).
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace dfchem {

constexpr double kGasConstant = 8314.46261815324;  // J/kmol/K (src_gpu/dfThermo.cu:9)
constexpr double kOneAtm = 101325.0;
constexpr double kCal = 4.184;
constexpr int kMaxSide = 3;  // max distinct species on one side of a reaction

enum ReactionKind : int32_t { kElementary = 0, kThreeBody = 1, kFalloffLindemann = 2, kFalloffTroe = 3 };

struct Mechanism {
    int N = 0;       // species
    int R = 0;       // reactions
    int n_elem = 0;  // elements
    std::vector<std::string> species;   // Cantera order = YAML phases[0].species order
    std::vector<std::string> elements;
    std::vector<double> comp;           // [N*n_elem] atoms of element e in species k
    std::vector<double> W;              // [N] kg/kmol
    std::vector<double> Tlo, Tmid, Thi; // [N]
    std::vector<double> nasa_lo;        // [N*7] range used for T <= Tmid
    std::vector<double> nasa_hi;        // [N*7]
    std::vector<double> hc;             // [N] Hf298SS/W, J/kg

    std::vector<int32_t> kind;          // [R] ReactionKind
    std::vector<int32_t> reversible;    // [R]
    std::vector<int32_t> r_idx, p_idx;  // [R*kMaxSide] species index or -1
    std::vector<double> A, b, Ta;       // [R] k = A T^b exp(-Ta/T) (high-pressure limit for falloff)
    std::vector<double> A0, b0, Ta0;    // [R] low-pressure limit (falloff), else 0
    std::vector<double> troe;           // [R*4] A, T3, T1, T2 (T2 = +inf when absent)
    std::vector<double> r_nu, r_ord;    // [R*kMaxSide] stoichiometric coefficient, forward order
    std::vector<double> p_nu;           // [R*kMaxSide]
    std::vector<double> eff;            // [R*N] third-body efficiencies (1 for M-less reactions)
    std::vector<std::string> equations; // [R] diagnostic only (not serialised)
    // Lennard-Jones / Stockmayer transport data of the species' `transport:` blocks (model: gas), [N*6]:
    // geometry (0 atom, 1 linear, 2 nonlinear), well-depth [K], diameter [Angstrom], dipole [Debye],
    // polarizability [Angstrom^3], rotational-relaxation.  Empty when any species lacks the block.
    std::vector<double> trans;

    // Flat little-endian blob "DFCHEMB1" (+ optional trailing section "TRN1" with `trans`); byte-identical to
    // oracle/mech.py::serialize().
    std::vector<uint8_t> serialize() const;
    static Mechanism deserialize(const uint8_t* data, size_t len);
};

// Parse a Cantera-YAML mechanism file. Throws std::runtime_error on anything outside the
// supported dialect (PLOG, Chebyshev, SRI, explicit colliders, non-NASA7 thermo, ...).
Mechanism compile_mechanism_file(const std::string& yaml_path);
Mechanism compile_mechanism_text(const std::string& yaml_text);

}  // namespace dfchem
