// io::parse_vasp -- minimal POSCAR (VASP 5) reader with the reference's conventions (src/io/vasp_parser.cpp:13-78):
// comment, scale, three lattice rows (multiplied by the scale), element symbols, counts, optional "Selective
// dynamics", Direct | Cartesian, one coordinate triple per atom; Cartesian input is converted to fractional with the
// inverse lattice.  Only used to load fixtures for the facade example -- the parser is outside the B200 path.
#include "io/vasp_parser.hpp"

#include <cctype>
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
    auto first = [](const std::string& s) {
        for (char ch : s)
            if (!std::isspace(static_cast<unsigned char>(ch))) return static_cast<char>(std::tolower(ch));
        return '\0';
    };
    if (first(mode) == 's') mode = next_line(in, filepath);  // Selective dynamics
    const bool cartesian = first(mode) == 'c' || first(mode) == 'k';
    long total = 0;
    for (size_t e = 0; e < out.counts.size(); ++e) {
        for (int k = 0; k < out.counts[e]; ++k) out.atom_types.push_back(static_cast<int>(e));
        total += out.counts[e];
    }
    out.frac_coords.resize(total, 3);
    const Eigen::Matrix3d inv_t = out.lattice.inverse().transpose();
    for (long i = 0; i < total; ++i) {
        std::istringstream ss(next_line(in, filepath));
        double a, b, c;
        if (!(ss >> a >> b >> c)) throw std::runtime_error("parse_vasp: bad coordinate line in " + filepath);
        Eigen::Vector3d p(a, b, c);
        if (cartesian) p = inv_t * (p * scale);
        out.frac_coords.row(i) = p;
    }
    return out;
}

}  // namespace defect_gnn::io
