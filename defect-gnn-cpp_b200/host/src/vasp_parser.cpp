// io::parse_vasp -- POSCAR (VASP 5) reader with the reference's semantics (src/io/vasp_parser.cpp:13-78; the rules
// that matter are spelled out in io/poscar_rules.hpp): comment, scale, three lattice rows (multiplied by the scale),
// element symbols, counts, Direct iff the mode line starts with 'd' / 'D', one coordinate triple per atom; Cartesian
// input is converted as cart * inverse(lattice) without scaling the coordinates, exactly like the reference.  Only used
// to load fixtures for the facade example -- the parser is outside the B200 path.
#include "io/vasp_parser.hpp"

#include "io/poscar_rules.hpp"

#include <fstream>
#include <sstream>
#include <stdexcept>

namespace defect_gnn::io {

namespace {
std::string next_line(std::ifstream& in, const std::string& path) {
    std::string line;
    if (!std::getline(in, line)) throw std::runtime_error("parse_vasp: unexpected end of file in " + path);
    return line;
}
}  // namespace

VASPStructure parse_vasp(const std::string& filepath) {
    std::ifstream in(filepath);
    if (!in) throw std::runtime_error("parse_vasp: cannot open " + filepath);
    VASPStructure out;
    next_line(in, filepath);  // comment
    double scale = 1.0;
    {
        std::istringstream ss(next_line(in, filepath));
        if (!(ss >> scale)) throw std::runtime_error("parse_vasp: bad scale in " + filepath);
    }
    for (int i = 0; i < 3; ++i) {
        std::istringstream ss(next_line(in, filepath));
        double a, b, c;
        if (!(ss >> a >> b >> c)) throw std::runtime_error("parse_vasp: bad lattice row in " + filepath);
        out.lattice(i, 0) = a;
        out.lattice(i, 1) = b;
        out.lattice(i, 2) = c;
    }
    out.lattice *= scale;
    {
        std::istringstream ss(next_line(in, filepath));
        std::string sym;
        while (ss >> sym) out.elements.push_back(sym);
    }
    {
        std::istringstream ss(next_line(in, filepath));
        int n;
        while (ss >> n) out.counts.push_back(n);
    }
    if (out.counts.empty() || out.counts.size() != out.elements.size())
        throw std::runtime_error("parse_vasp: element / count lines do not match in " + filepath);
    std::string mode = next_line(in, filepath);
    if (rules::is_selective(mode.data(), mode.data() + mode.size())) mode = next_line(in, filepath);  // (see poscar_rules.hpp)
    const bool direct = rules::is_direct(mode.data(), mode.data() + mode.size());
    long total = 0;
    for (size_t e = 0; e < out.counts.size(); ++e) {
        for (int k = 0; k < out.counts[e]; ++k) out.atom_types.push_back(static_cast<int>(e));
        total += out.counts[e];
    }
    out.frac_coords.resize(total, 3);
    double L[9], inv[9];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) L[3 * i + j] = out.lattice(i, j);
    if (!direct) rules::inverse3(L, inv);
    for (long i = 0; i < total; ++i) {
        std::istringstream ss(next_line(in, filepath));
        double c[3];
        if (!(ss >> c[0] >> c[1] >> c[2])) throw std::runtime_error("parse_vasp: bad coordinate line in " + filepath);
        if (!direct) {
            double f[3];
            rules::cart_to_frac_row(i, total, c, inv, f);
            c[0] = f[0]; c[1] = f[1]; c[2] = f[2];
        }
        out.frac_coords.row(i) = Eigen::Vector3d(c[0], c[1], c[2]);
    }
    return out;
}

}  // namespace defect_gnn::io
