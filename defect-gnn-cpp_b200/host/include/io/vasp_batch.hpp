// io/vasp_batch.hpp -- NEW surface: bulk POSCAR -> SoA ingest.  At > 10^6 structures/s per GPU the
// reference's one-file-at-a-time std::stringstream parser (src/io/vasp_parser.cpp:13-78; string variant
// src/viz/wasm_bindings.cpp:91-116) is what starves the device, so files are parsed by a pool of threads with a
// strtod tokenizer straight into the SoA buffers dgnn_build reads (graph::SoABatch).  Parser semantics are the
// reference's (io/poscar_rules.hpp): line 2 scale, lines 3-5 lattice rows (* scale), line 6 symbols, line 7 counts,
// line 8 Direct iff it starts with 'd' / 'D', one triple per atom; Cartesian triples are converted as
// cart * inverse(lattice), unscaled, like vasp_parser.cpp:72-75;
// positions = lattice^T * frac exactly as crystal::Structure computes them (src/crystal/structure.cpp:17).
#pragma once

#include "graph/batch.hpp"

#include <string>
#include <vector>

namespace defect_gnn::io {

// One POSCAR given as text appended to the batch.  Throws std::runtime_error on malformed input.
void append_vasp_text(const char* text, size_t len, graph::SoABatch& batch);

// Files parsed in parallel (n_threads = 0: hardware concurrency), appended in the order of `paths`.
graph::SoABatch load_vasp_batch(const std::vector<std::string>& paths, unsigned n_threads = 0);

// The same ingest into PAGE-LOCKED SoA buffers (graph::PinnedSoABatch): headers first (atom counts -> offsets), then the
// coordinate lines are parsed straight into their final place in pinned memory, which BatchBuilder::build hands to
// dgnn_build's chunked upload queue as is.  Needs the CUDA driver (pinned allocation): throws without a device.
graph::PinnedSoABatch load_vasp_batch_pinned(const std::vector<std::string>& paths, unsigned n_threads = 0);

}  // namespace defect_gnn::io
