// Transport-fit generator (transport_fit.cpp): visc[N][5], cond[N][5], diff[N][N][5] in the layout of the
// reference's thermo_<mech>.txt (src_gpu/dfThermo.cu:361-435).  Returns false when the mechanism carries no
// (complete) `transport:` data.
#pragma once
#include <vector>

#include "mechanism.hpp"

namespace dfchem {
bool generate_transport_fits(const Mechanism& m, std::vector<double>& visc, std::vector<double>& cond, std::vector<double>& diff);
}
