// io::load_vasp_batch -- multi-threaded POSCAR -> SoA ingest (see io/vasp_batch.hpp).
#include "io/vasp_batch.hpp"

#include "io/poscar_rules.hpp"

#include <atomic>
#include <cctype>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <thread>

namespace defect_gnn::io {

namespace {

struct Cursor {
    const char* p;
    const char* end;
    // [b, e) of the next line (without the terminator); false at end of input
    bool line(const char*& b, const char*& e) {
        if (p >= end) return false;
        b = p;
        const char* nl = static_cast<const char*>(memchr(p, '\n', static_cast<size_t>(end - p)));
        e = nl ? nl : end;
        p = nl ? nl + 1 : end;
        if (e > b && e[-1] == '\r') --e;
        return true;
    }
};

// up to `want` doubles from [b, e); returns how many were read (operator>> stops at the first non-number, too)
int read_doubles(const char* b, const char* e, double* out, int want) {
    char buf[512];
    const size_t n = static_cast<size_t>(e - b) < sizeof(buf) - 1 ? static_cast<size_t>(e - b) : sizeof(buf) - 1;
    memcpy(buf, b, n);
    buf[n] = '\0';
    char* q = buf;
    int got = 0;
    while (got < want) {
        char* r = nullptr;
        const double v = strtod(q, &r);
        if (r == q) break;
        out[got++] = v;
        q = r;
    }
    return got;
}

struct Parsed {
    double lattice[9];
    std::vector<double> x, y, z;
    std::vector<int> element;
};

// everything before the coordinate lines; `rest` is left at the first of them
struct Header {
    double lattice[9];
    double inv[9];
    std::vector<int> counts;
    long total = 0;
    bool cartesian = false;
    Cursor rest{nullptr, nullptr};
};

void parse_header(const char* text, size_t len, Header& h, const std::string& what) {
    Cursor c{text, text + len};
    const char *b, *e;
    auto need = [&](const char* which) {
        if (!c.line(b, e)) throw std::runtime_error("parse_vasp: unexpected end of input (" + std::string(which) + ") in " + what);
    };
    need("comment");
    need("scale");
    double scale = 1.0;
    if (read_doubles(b, e, &scale, 1) != 1) throw std::runtime_error("parse_vasp: bad scale in " + what);
    for (int i = 0; i < 3; ++i) {
        need("lattice");
        if (read_doubles(b, e, h.lattice + 3 * i, 3) != 3) throw std::runtime_error("parse_vasp: bad lattice row in " + what);
    }
    for (double& v : h.lattice) v *= scale;  // vasp.lattice *= scale_factor
    need("elements");
    int n_sym = 0;
    for (const char* q = b; q < e;) {
        while (q < e && isspace(static_cast<unsigned char>(*q))) ++q;
        if (q < e) ++n_sym;
        while (q < e && !isspace(static_cast<unsigned char>(*q))) ++q;
    }
    need("counts");
    {
        double tmp[64];
        const int n = read_doubles(b, e, tmp, 64);
        for (int i = 0; i < n; ++i) h.counts.push_back(static_cast<int>(tmp[i]));
    }
    if (h.counts.empty() || static_cast<int>(h.counts.size()) != n_sym)
        throw std::runtime_error("parse_vasp: element / count lines do not match in " + what);
    need("mode");
    if (rules::is_selective(b, e)) need("mode");  // (see io/poscar_rules.hpp)
    h.cartesian = !rules::is_direct(b, e);
    for (double& v : h.inv) v = 0.0;
    if (h.cartesian) rules::inverse3(h.lattice, h.inv);
    h.total = 0;
    for (int n : h.counts) h.total += n < 0 ? 0 : n;
    h.rest = c;
}

// the coordinate lines -> Cartesian positions written straight to their destination (h.total entries each)
void parse_atoms(Header& h, double* x, double* y, double* z, int* element, const std::string& what) {
    const double* L = h.lattice;
    const char *b, *e;
    long i = 0;
    for (size_t el = 0; el < h.counts.size(); ++el) {
        for (int k = 0; k < h.counts[el]; ++k, ++i) {
            if (!h.rest.line(b, e)) throw std::runtime_error("parse_vasp: unexpected end of input (coordinates) in " + what);
            double f[3];
            if (read_doubles(b, e, f, 3) != 3) throw std::runtime_error("parse_vasp: bad coordinate line in " + what);
            if (h.cartesian) {  // frac = cart * inverse(lattice), coordinates unscaled (vasp_parser.cpp:72-75)
                double g[3];
                rules::cart_to_frac_row(i, h.total, f, h.inv, g);
                f[0] = g[0]; f[1] = g[1]; f[2] = g[2];
            }
            // position = lattice^T * frac (structure.cpp:17): component k = (e0 + e1) + e2, e_r = L(r,k) f_r -- the
            // order of Eigen 3.4's vectorised 3-term reduction on x86-64 (pinned on oracle/_ref; same as the facade's Structure)
            x[i] = (L[0] * f[0] + L[3] * f[1]) + L[6] * f[2];
            y[i] = (L[1] * f[0] + L[4] * f[1]) + L[7] * f[2];
            z[i] = (L[2] * f[0] + L[5] * f[1]) + L[8] * f[2];
            element[i] = static_cast<int>(el);
        }
    }
}

void parse_text(const char* text, size_t len, Parsed& out, const std::string& what) {
    Header h;
    parse_header(text, len, h, what);
    for (int k = 0; k < 9; ++k) out.lattice[k] = h.lattice[k];
    const size_t n = static_cast<size_t>(h.total);
    out.x.resize(n);
    out.y.resize(n);
    out.z.resize(n);
    out.element.resize(n);
    parse_atoms(h, out.x.data(), out.y.data(), out.z.data(), out.element.data(), what);
}

std::string read_file(const std::string& path) {
    FILE* f = fopen(path.c_str(), "rb");
    if (!f) throw std::runtime_error("parse_vasp: cannot open " + path);
    std::string buf;
    char chunk[1 << 16];
    size_t got;
    while ((got = fread(chunk, 1, sizeof(chunk), f)) > 0) buf.append(chunk, got);
    fclose(f);
    return buf;
}

template <class F>
void run_parallel(size_t n, unsigned n_threads, F&& body) {
    if (n_threads == 0) n_threads = std::thread::hardware_concurrency();
    if (n_threads == 0) n_threads = 1;
    if (n_threads > n) n_threads = static_cast<unsigned>(n > 0 ? n : 1);
    std::atomic<size_t> next{0};
    auto work = [&]() {
        for (size_t i = next.fetch_add(1); i < n; i = next.fetch_add(1)) body(i);
    };
    std::vector<std::thread> pool;
    for (unsigned t = 1; t < n_threads; ++t) pool.emplace_back(work);
    work();
    for (auto& t : pool) t.join();
}

void append(const Parsed& p, graph::SoABatch& batch) {
    batch.x.insert(batch.x.end(), p.x.begin(), p.x.end());
    batch.y.insert(batch.y.end(), p.y.begin(), p.y.end());
    batch.z.insert(batch.z.end(), p.z.begin(), p.z.end());
    batch.element.insert(batch.element.end(), p.element.begin(), p.element.end());
    batch.lattice.insert(batch.lattice.end(), p.lattice, p.lattice + 9);
    batch.atom_offsets.push_back(batch.atom_offsets.back() + static_cast<int64_t>(p.x.size()));
}

}  // namespace

void append_vasp_text(const char* text, size_t len, graph::SoABatch& batch) {
    Parsed p;
    parse_text(text, len, p, "<text>");
    append(p, batch);
}

graph::SoABatch load_vasp_batch(const std::vector<std::string>& paths, unsigned n_threads) {
    const size_t n = paths.size();
    std::vector<Parsed> parsed(n);
    std::vector<std::string> errors(n);
    run_parallel(n, n_threads, [&](size_t i) {
        try {
            const std::string buf = read_file(paths[i]);
            parse_text(buf.data(), buf.size(), parsed[i], paths[i]);
        } catch (const std::exception& ex) {
            errors[i] = ex.what();
        }
    });
    for (size_t i = 0; i < n; ++i)
        if (!errors[i].empty()) throw std::runtime_error(errors[i]);
    graph::SoABatch batch;
    size_t total = 0;
    for (const Parsed& p : parsed) total += p.x.size();
    batch.x.reserve(total);
    batch.y.reserve(total);
    batch.z.reserve(total);
    batch.element.reserve(total);
    batch.lattice.reserve(9 * n);
    batch.atom_offsets.reserve(n + 1);
    for (const Parsed& p : parsed) append(p, batch);
    return batch;
}

// Pinned ingest: pass 1 (parallel) reads every file and parses its header -- the atom count is on line 7 -- , a prefix
// sum places the structures, the pinned SoA buffers are allocated once, and pass 2 (parallel) parses the coordinate
// lines STRAIGHT into their final place in page-locked memory: what dgnn_build's copy queue uploads, no staging copy.
graph::PinnedSoABatch load_vasp_batch_pinned(const std::vector<std::string>& paths, unsigned n_threads) {
    const size_t n = paths.size();
    std::vector<std::string> text(n), errors(n);
    std::vector<Header> hdr(n);
    run_parallel(n, n_threads, [&](size_t i) {
        try {
            text[i] = read_file(paths[i]);
            parse_header(text[i].data(), text[i].size(), hdr[i], paths[i]);
        } catch (const std::exception& ex) {
            errors[i] = ex.what();
        }
    });
    for (size_t i = 0; i < n; ++i)
        if (!errors[i].empty()) throw std::runtime_error(errors[i]);
    int64_t total = 0;
    for (const Header& h : hdr) total += h.total;
    graph::PinnedSoABatch batch(static_cast<int64_t>(n), total);
    batch.atom_offsets[0] = 0;
    for (size_t i = 0; i < n; ++i) batch.atom_offsets[i + 1] = batch.atom_offsets[i] + hdr[i].total;
    run_parallel(n, n_threads, [&](size_t i) {
        try {
            const int64_t a0 = batch.atom_offsets[i];
            for (int k = 0; k < 9; ++k) batch.lattice[9 * i + k] = hdr[i].lattice[k];
            parse_atoms(hdr[i], batch.x + a0, batch.y + a0, batch.z + a0, batch.element + a0, paths[i]);
        } catch (const std::exception& ex) {
            errors[i] = ex.what();
        }
    });
    for (size_t i = 0; i < n; ++i)
        if (!errors[i].empty()) throw std::runtime_error(errors[i]);
    return batch;
}

}  // namespace defect_gnn::io
