// Host mechanism compiler: Cantera-YAML mechanism file -> flat Mechanism record.
//
// Replaces, for the hot path only, what the reference gets from Cantera::newSolution
// (src/dfCanteraMixture/CanteraMixture.C:65): species list and order, molecular weights,
// NASA-7 tables, Hf298 (for hc_, src/dfChemistryModel/dfChemistryModel.C:335-338) and the
// reaction set (Arrhenius / three-body / Lindemann+Troe falloff, `orders` overrides,
// efficiencies, duplicates) of mechanisms/{H2,CH4,C7H16,C2H5OH}.  Unit handling follows the
// Cantera YAML rules stated in SURVEY.md Appendix A.1.
#include <cmath>
#include <cstring>
#include <limits>
#include <map>
#include <stdexcept>

#include "mechanism.hpp"
#include "yaml_lite.hpp"

namespace dfchem {
namespace {

// Cantera 2.6 Elements.cpp values for the elements appearing in mechanisms/
double atomic_weight(const std::string& e) {
    static const std::map<std::string, double> tab = {
        {"H", 1.008}, {"He", 4.002602}, {"HE", 4.002602}, {"C", 12.011}, {"N", 14.007},
        {"O", 15.999}, {"Ar", 39.948},  {"AR", 39.948},
    };
    auto it = tab.find(e);
    if (it == tab.end()) throw std::runtime_error("mechanism: unknown element '" + e + "'");
    return it->second;
}

double unit_factor(const std::map<std::string, double>& tab, const std::string& u, const char* what) {
    auto it = tab.find(u);
    if (it == tab.end()) throw std::runtime_error(std::string("mechanism: unsupported ") + what + " unit '" + u + "'");
    return it->second;
}

const std::map<std::string, double> kLen = {{"m", 1.0}, {"cm", 1e-2}, {"mm", 1e-3}};
const std::map<std::string, double> kQty = {{"kmol", 1.0}, {"mol", 1e-3}, {"molec", 1.0 / 6.02214076e26}};
const std::map<std::string, double> kTime = {{"s", 1.0}, {"ms", 1e-3}, {"min", 60.0}};
const std::map<std::string, double> kEa = {  // -> J/kmol
    {"J/kmol", 1.0},          {"J/mol", 1e3},          {"kJ/mol", 1e6},
    {"cal/mol", kCal * 1e3},  {"kcal/mol", kCal * 1e6}, {"K", kGasConstant}};

// "1.033e+05 cal/mol" -> value + unit ("" when bare)
void num_unit(const yaml::Node& n, double& v, std::string& unit) {
    const std::string& s = n.as_string();
    size_t sp = s.find(' ');
    if (sp == std::string::npos) {
        v = n.as_double();
        unit.clear();
        return;
    }
    v = std::strtod(s.substr(0, sp).c_str(), nullptr);
    size_t a = sp;
    while (a < s.size() && s[a] == ' ') ++a;
    unit = s.substr(a);
}

struct Term {
    std::string name;
    double nu;
};

std::vector<std::string> split(const std::string& s, const std::string& sep) {
    std::vector<std::string> out;
    size_t a = 0;
    while (true) {
        size_t b = s.find(sep, a);
        if (b == std::string::npos) {
            out.push_back(s.substr(a));
            break;
        }
        out.push_back(s.substr(a, b - a));
        a = b + sep.size();
    }
    return out;
}
std::string trim(const std::string& s) {
    size_t a = 0, b = s.size();
    while (a < b && s[a] == ' ') ++a;
    while (b > a && s[b - 1] == ' ') --b;
    return s.substr(a, b - a);
}

// one side of an equation: "2 H + M" -> [(H,2)], has_M
void parse_side(const std::string& text, std::vector<Term>& out, bool& has_m) {
    for (const std::string& raw : split(text, " + ")) {
        std::string tok = trim(raw);
        if (tok.empty()) continue;
        double nu = 1.0;
        std::string name = tok;
        size_t sp = tok.find(' ');
        if (sp != std::string::npos) {
            char* end = nullptr;
            double v = std::strtod(tok.c_str(), &end);
            if (end == tok.c_str() + sp) {
                nu = v;
                name = trim(tok.substr(sp));
            }
        }
        if (name == "M") {
            has_m = true;
            continue;
        }
        bool merged = false;
        for (auto& t : out)
            if (t.name == name) {
                t.nu += nu;
                merged = true;
            }
        if (!merged) out.push_back({name, nu});
    }
}

struct Equation {
    std::vector<Term> reactants, products;
    bool reversible = true, has_m = false, falloff = false;
};

Equation parse_equation(std::string eq) {
    Equation e;
    // remove "(+M)" / "(+ M)"
    for (const char* pat : {"(+M)", "(+ M)"}) {
        size_t p;
        while ((p = eq.find(pat)) != std::string::npos) {
            eq.erase(p, std::strlen(pat));
            e.falloff = true;
        }
    }
    if (eq.find("(+") != std::string::npos)
        throw std::runtime_error("mechanism: explicit falloff collider not supported: " + eq);
    std::string lhs, rhs;
    size_t p;
    if ((p = eq.find("<=>")) != std::string::npos) {
        lhs = eq.substr(0, p);
        rhs = eq.substr(p + 3);
        e.reversible = true;
    } else if ((p = eq.find("=>")) != std::string::npos) {
        lhs = eq.substr(0, p);
        rhs = eq.substr(p + 2);
        e.reversible = false;
    } else if ((p = eq.find('=')) != std::string::npos) {
        lhs = eq.substr(0, p);
        rhs = eq.substr(p + 1);
        e.reversible = true;
    } else {
        throw std::runtime_error("mechanism: cannot parse equation '" + eq + "'");
    }
    bool m1 = false, m2 = false;
    parse_side(lhs, e.reactants, m1);
    parse_side(rhs, e.products, m2);
    e.has_m = m1 || m2;
    return e;
}

double nasa_h_RT(const double* a, double T) {
    return a[0] + a[1] * T / 2 + a[2] * std::pow(T, 2.0) / 3 + a[3] * std::pow(T, 3.0) / 4 +
           a[4] * std::pow(T, 4.0) / 5 + a[5] / T;
}

Mechanism compile(const yaml::Node& doc) {
    Mechanism m;
    double u_len = 1.0, u_qty = 1.0, u_time = 1.0, u_ea = 1.0;
    if (doc.has("units") && doc.at("units").is_map()) {
        const yaml::Node& u = doc.at("units");
        if (u.has("length")) u_len = unit_factor(kLen, u.at("length").as_string(), "length");
        if (u.has("quantity")) u_qty = unit_factor(kQty, u.at("quantity").as_string(), "quantity");
        if (u.has("time")) u_time = unit_factor(kTime, u.at("time").as_string(), "time");
        if (u.has("activation-energy")) {
            u_ea = unit_factor(kEa, u.at("activation-energy").as_string(), "activation-energy");
        } else {
            u_ea = u.has("quantity") ? 1.0 / u_qty : 1.0;
            if (u.has("energy")) {
                static const std::map<std::string, double> en = {{"J", 1.0}, {"cal", kCal}, {"kJ", 1e3}, {"kcal", kCal * 1e3}};
                u_ea *= unit_factor(en, u.at("energy").as_string(), "energy");
            }
        }
    }
    const double conc = std::pow(u_len, 3.0) / u_qty;  // (length^3/quantity) in m^3/kmol

    const yaml::Node& phase = doc.at("phases").at(0);
    const yaml::Node& sp_defs = doc.at("species");
    std::map<std::string, const yaml::Node*> by_name;
    for (size_t i = 0; i < sp_defs.size(); ++i) by_name[sp_defs.at(i).at("name").as_string()] = &sp_defs.at(i);

    const yaml::Node& psp = phase.at("species");
    if (psp.is_seq() && !(psp.size() == 1 && psp.at(0).is_scalar() && psp.at(0).as_string() == "all")) {
        for (size_t i = 0; i < psp.size(); ++i) m.species.push_back(psp.at(i).as_string());
    } else {
        for (size_t i = 0; i < sp_defs.size(); ++i) m.species.push_back(sp_defs.at(i).at("name").as_string());
    }
    if (phase.has("elements"))
        for (size_t i = 0; i < phase.at("elements").size(); ++i) m.elements.push_back(phase.at("elements").at(i).as_string());
    if (phase.has("thermo") && phase.at("thermo").as_string() != "ideal-gas")
        throw std::runtime_error("mechanism: only thermo: ideal-gas is supported (real-fluid phases are out of scope)");
    m.N = (int)m.species.size();
    m.n_elem = (int)m.elements.size();
    const int N = m.N;
    auto sp_index = [&](const std::string& n) -> int {
        for (int k = 0; k < N; ++k)
            if (m.species[k] == n) return k;
        return -1;
    };
    m.comp.assign((size_t)N * m.n_elem, 0.0);
    m.W.assign(N, 0.0);
    m.Tlo.assign(N, 0.0);
    m.Tmid.assign(N, 0.0);
    m.Thi.assign(N, 0.0);
    m.nasa_lo.assign((size_t)N * 7, 0.0);
    m.nasa_hi.assign((size_t)N * 7, 0.0);
    m.hc.assign(N, 0.0);
    for (int k = 0; k < N; ++k) {
        auto it = by_name.find(m.species[k]);
        if (it == by_name.end()) throw std::runtime_error("mechanism: species '" + m.species[k] + "' not defined");
        const yaml::Node& s = *it->second;
        for (auto& kv : s.at("composition").map) {
            int e = -1;
            for (int j = 0; j < m.n_elem; ++j)
                if (m.elements[j] == kv.first) e = j;
            if (e < 0) throw std::runtime_error("mechanism: element '" + kv.first + "' not in phase");
            double c = kv.second->as_double();
            m.comp[(size_t)k * m.n_elem + e] = c;
            m.W[k] += atomic_weight(kv.first) * c;
        }
        const yaml::Node& th = s.at("thermo");
        if (th.at("model").as_string() != "NASA7") throw std::runtime_error("mechanism: only NASA7 thermo is supported");
        const yaml::Node& tr = th.at("temperature-ranges");
        const yaml::Node& data = th.at("data");
        auto row = [&](size_t r, double* dst) {
            if (data.at(r).size() != 7) throw std::runtime_error("mechanism: NASA7 row needs 7 coefficients");
            for (int j = 0; j < 7; ++j) dst[j] = data.at(r).at(j).as_double();
        };
        if (tr.size() == 2) {  // single range
            m.Tlo[k] = tr.at(0).as_double();
            m.Tmid[k] = tr.at(1).as_double();
            m.Thi[k] = tr.at(1).as_double();
            row(0, &m.nasa_lo[(size_t)k * 7]);
            row(0, &m.nasa_hi[(size_t)k * 7]);
        } else if (tr.size() == 3) {
            m.Tlo[k] = tr.at(0).as_double();
            m.Tmid[k] = tr.at(1).as_double();
            m.Thi[k] = tr.at(2).as_double();
            row(0, &m.nasa_lo[(size_t)k * 7]);
            row(1, &m.nasa_hi[(size_t)k * 7]);
        } else {
            throw std::runtime_error("mechanism: NASA7 needs 1 or 2 temperature ranges");
        }
        const double T0 = 298.15;
        const double* a = T0 <= m.Tmid[k] ? &m.nasa_lo[(size_t)k * 7] : &m.nasa_hi[(size_t)k * 7];
        m.hc[k] = (nasa_h_RT(a, T0) * kGasConstant * T0) / m.W[k];
    }
    // transport data (input of the fit generator, transport_fit.cpp); all species or none
    m.trans.assign((size_t)N * 6, 0.0);
    for (int k = 0; k < N && !m.trans.empty(); ++k) {
        const yaml::Node& s = *by_name.find(m.species[k])->second;
        if (!s.has("transport") || !s.at("transport").has("geometry") || !s.at("transport").has("well-depth") ||
            !s.at("transport").has("diameter")) {
            m.trans.clear();
            break;
        }
        const yaml::Node& t = s.at("transport");
        const std::string g = t.at("geometry").as_string();
        double* r = &m.trans[(size_t)k * 6];
        r[0] = g == "atom" ? 0.0 : (g == "linear" ? 1.0 : 2.0);
        r[1] = t.at("well-depth").as_double();
        r[2] = t.at("diameter").as_double();
        r[3] = t.has("dipole") ? t.at("dipole").as_double() : 0.0;
        r[4] = t.has("polarizability") ? t.at("polarizability").as_double() : 0.0;
        r[5] = t.has("rotational-relaxation") ? t.at("rotational-relaxation").as_double() : 0.0;
    }

    const yaml::Node& rx = doc.at("reactions");
    const int R = m.R = (int)rx.size();
    m.kind.assign(R, 0);
    m.reversible.assign(R, 0);
    m.r_idx.assign((size_t)R * kMaxSide, -1);
    m.p_idx.assign((size_t)R * kMaxSide, -1);
    for (auto* v : {&m.A, &m.b, &m.Ta, &m.A0, &m.b0, &m.Ta0}) v->assign(R, 0.0);
    m.troe.assign((size_t)R * 4, 0.0);
    m.r_nu.assign((size_t)R * kMaxSide, 0.0);
    m.r_ord.assign((size_t)R * kMaxSide, 0.0);
    m.p_nu.assign((size_t)R * kMaxSide, 0.0);
    m.eff.assign((size_t)R * N, 1.0);
    m.equations.resize(R);

    auto arrhenius = [&](const yaml::Node& rc, double order, double& A, double& b, double& Ta) {
        double a, e;
        std::string ua, ue;
        num_unit(rc.at("A"), a, ua);
        if (!ua.empty()) throw std::runtime_error("mechanism: explicit units on A are not supported");
        num_unit(rc.at("Ea"), e, ue);
        double ea = e * (ue.empty() ? u_ea : unit_factor(kEa, ue, "activation-energy"));
        A = a * std::pow(conc, order - 1) / u_time;
        b = rc.at("b").as_double();
        Ta = ea / kGasConstant;
    };

    for (int i = 0; i < R; ++i) {
        const yaml::Node& r = rx.at(i);
        for (const char* bad : {"negative-A", "rate-constants", "data", "SRI", "Tsang", "coverage-dependencies"})
            if (r.has(bad)) throw std::runtime_error(std::string("mechanism: unsupported reaction field '") + bad + "'");
        m.equations[i] = r.at("equation").as_string();
        Equation e = parse_equation(m.equations[i]);
        std::string typ = r.has("type") ? r.at("type").as_string() : "elementary";
        if ((int)e.reactants.size() > kMaxSide || (int)e.products.size() > kMaxSide)
            throw std::runtime_error("mechanism: more than 3 distinct species on one side: " + m.equations[i]);
        double order = 0.0;
        for (size_t j = 0; j < e.reactants.size(); ++j) {
            int k = sp_index(e.reactants[j].name);
            if (k < 0) throw std::runtime_error("mechanism: unknown species '" + e.reactants[j].name + "' in " + m.equations[i]);
            m.r_idx[(size_t)i * kMaxSide + j] = k;
            m.r_nu[(size_t)i * kMaxSide + j] = e.reactants[j].nu;
            double ord = e.reactants[j].nu;
            if (r.has("orders") && r.at("orders").has(e.reactants[j].name)) ord = r.at("orders").at(e.reactants[j].name).as_double();
            m.r_ord[(size_t)i * kMaxSide + j] = ord;
        }
        if (r.has("orders")) {
            if (e.reversible) throw std::runtime_error("mechanism: 'orders' requires an irreversible reaction: " + m.equations[i]);
            for (auto& kv : r.at("orders").map) {
                bool found = false;
                for (auto& t : e.reactants) found |= t.name == kv.first;
                if (!found) throw std::runtime_error("mechanism: order given for non-reactant '" + kv.first + "'");
            }
        }
        for (int j = 0; j < kMaxSide; ++j) order += m.r_ord[(size_t)i * kMaxSide + j];
        for (size_t j = 0; j < e.products.size(); ++j) {
            int k = sp_index(e.products[j].name);
            if (k < 0) throw std::runtime_error("mechanism: unknown species '" + e.products[j].name + "' in " + m.equations[i]);
            m.p_idx[(size_t)i * kMaxSide + j] = k;
            m.p_nu[(size_t)i * kMaxSide + j] = e.products[j].nu;
        }
        m.reversible[i] = e.reversible ? 1 : 0;
        if (typ == "falloff") {
            if (!e.falloff) throw std::runtime_error("mechanism: falloff reaction without (+M): " + m.equations[i]);
            m.kind[i] = r.has("Troe") ? kFalloffTroe : kFalloffLindemann;
            arrhenius(r.at("high-P-rate-constant"), order, m.A[i], m.b[i], m.Ta[i]);
            arrhenius(r.at("low-P-rate-constant"), order + 1, m.A0[i], m.b0[i], m.Ta0[i]);
            if (r.has("Troe")) {
                const yaml::Node& t = r.at("Troe");
                m.troe[(size_t)i * 4 + 0] = t.at("A").as_double();
                m.troe[(size_t)i * 4 + 1] = t.at("T3").as_double();
                m.troe[(size_t)i * 4 + 2] = t.at("T1").as_double();
                m.troe[(size_t)i * 4 + 3] = t.has("T2") ? t.at("T2").as_double() : std::numeric_limits<double>::infinity();
                // Cantera 2.6 (TroeRate::setFalloffCoeffs): T2 = 0 means "no exp(-T2/T) term", not exp(0) = 1
                if (m.troe[(size_t)i * 4 + 3] == 0.0) m.troe[(size_t)i * 4 + 3] = std::numeric_limits<double>::infinity();
            }
        } else if (typ == "three-body") {
            if (!e.has_m) throw std::runtime_error("mechanism: three-body reaction without M: " + m.equations[i]);
            m.kind[i] = kThreeBody;
            arrhenius(r.at("rate-constant"), order + 1, m.A[i], m.b[i], m.Ta[i]);
        } else if (typ == "elementary") {
            if (e.has_m || e.falloff) throw std::runtime_error("mechanism: M in an elementary reaction: " + m.equations[i]);
            m.kind[i] = kElementary;
            arrhenius(r.at("rate-constant"), order, m.A[i], m.b[i], m.Ta[i]);
        } else {
            throw std::runtime_error("mechanism: unsupported reaction type '" + typ + "'");
        }
        if (m.kind[i] != kElementary) {
            double def = r.has("default-efficiency") ? r.at("default-efficiency").as_double() : 1.0;
            for (int k = 0; k < N; ++k) m.eff[(size_t)i * N + k] = def;
            if (r.has("efficiencies"))
                for (auto& kv : r.at("efficiencies").map) {
                    int k = sp_index(kv.first);
                    if (k >= 0) m.eff[(size_t)i * N + k] = kv.second->as_double();
                }
        }
    }
    return m;
}

template <class T>
void put(std::vector<uint8_t>& out, const std::vector<T>& v) {
    const uint8_t* p = reinterpret_cast<const uint8_t*>(v.data());
    out.insert(out.end(), p, p + v.size() * sizeof(T));
}
template <class T>
void get(const uint8_t*& p, const uint8_t* end, std::vector<T>& v, size_t n) {
    if ((size_t)(end - p) < n * sizeof(T)) throw std::runtime_error("mechanism blob: truncated");
    v.resize(n);
    std::memcpy(v.data(), p, n * sizeof(T));
    p += n * sizeof(T);
}

}  // namespace

std::vector<uint8_t> Mechanism::serialize() const {
    std::vector<uint8_t> out;
    const char magic[8] = {'D', 'F', 'C', 'H', 'E', 'M', 'B', '1'};
    out.insert(out.end(), magic, magic + 8);
    int32_t hdr[4] = {N, R, n_elem, kMaxSide};
    out.insert(out.end(), reinterpret_cast<uint8_t*>(hdr), reinterpret_cast<uint8_t*>(hdr) + sizeof(hdr));
    for (auto& s : species) {
        char buf[32] = {0};
        std::memcpy(buf, s.c_str(), s.size() < 32 ? s.size() : 32);
        out.insert(out.end(), buf, buf + 32);
    }
    for (auto& s : elements) {
        char buf[8] = {0};
        std::memcpy(buf, s.c_str(), s.size() < 8 ? s.size() : 8);
        out.insert(out.end(), buf, buf + 8);
    }
    put(out, comp); put(out, W); put(out, Tlo); put(out, Tmid); put(out, Thi);
    put(out, nasa_lo); put(out, nasa_hi); put(out, hc);
    put(out, kind); put(out, reversible); put(out, r_idx); put(out, p_idx);
    put(out, A); put(out, b); put(out, Ta); put(out, A0); put(out, b0); put(out, Ta0);
    put(out, troe); put(out, r_nu); put(out, r_ord); put(out, p_nu); put(out, eff);
    if (!trans.empty()) {
        const char sec[8] = {'T', 'R', 'N', '1', 0, 0, 0, 0};
        out.insert(out.end(), sec, sec + 8);
        put(out, trans);
    }
    return out;
}

Mechanism Mechanism::deserialize(const uint8_t* data, size_t len) {
    if (len < 24 || std::memcmp(data, "DFCHEMB1", 8) != 0) throw std::runtime_error("mechanism blob: bad magic");
    Mechanism m;
    int32_t hdr[4];
    std::memcpy(hdr, data + 8, 16);
    m.N = hdr[0]; m.R = hdr[1]; m.n_elem = hdr[2];
    if (hdr[3] != kMaxSide || m.N <= 0 || m.R < 0) throw std::runtime_error("mechanism blob: bad header");
    const uint8_t* p = data + 24;
    const uint8_t* end = data + len;
    const size_t N = m.N, R = m.R, E = m.n_elem;
    if ((size_t)(end - p) < 32 * N + 8 * E) throw std::runtime_error("mechanism blob: truncated");
    for (size_t k = 0; k < N; ++k, p += 32) m.species.emplace_back(reinterpret_cast<const char*>(p), strnlen(reinterpret_cast<const char*>(p), 32));
    for (size_t k = 0; k < E; ++k, p += 8) m.elements.emplace_back(reinterpret_cast<const char*>(p), strnlen(reinterpret_cast<const char*>(p), 8));
    get(p, end, m.comp, N * E); get(p, end, m.W, N); get(p, end, m.Tlo, N); get(p, end, m.Tmid, N); get(p, end, m.Thi, N);
    get(p, end, m.nasa_lo, N * 7); get(p, end, m.nasa_hi, N * 7); get(p, end, m.hc, N);
    get(p, end, m.kind, R); get(p, end, m.reversible, R); get(p, end, m.r_idx, R * kMaxSide); get(p, end, m.p_idx, R * kMaxSide);
    get(p, end, m.A, R); get(p, end, m.b, R); get(p, end, m.Ta, R); get(p, end, m.A0, R); get(p, end, m.b0, R); get(p, end, m.Ta0, R);
    get(p, end, m.troe, R * 4); get(p, end, m.r_nu, R * kMaxSide); get(p, end, m.r_ord, R * kMaxSide); get(p, end, m.p_nu, R * kMaxSide);
    get(p, end, m.eff, R * N);
    if ((size_t)(end - p) >= 8 && std::memcmp(p, "TRN1", 4) == 0) {
        p += 8;
        get(p, end, m.trans, N * 6);
    }
    if (p != end) throw std::runtime_error("mechanism blob: trailing bytes");
    m.equations.assign(R, "");
    return m;
}

Mechanism compile_mechanism_text(const std::string& yaml_text) {
    yaml::NodePtr doc = yaml::parse(yaml_text);
    return compile(*doc);
}

Mechanism compile_mechanism_file(const std::string& yaml_path) {
    yaml::NodePtr doc = yaml::parse_file(yaml_path);
    return compile(*doc);
}

}  // namespace dfchem
