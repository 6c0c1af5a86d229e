// Host-side circuit "pass compiler" for the large-state engine (K5/K6): turns a gate list into a
// short sequence of passes, each of which touches a set of at most `tile_bits` qubits so that a
// CTA can pull 2^tile_bits amplitudes into shared memory, apply every gate of the pass there, and
// write them back - several gates per trip through HBM/L2 instead of one.
//
// Scheduling is a dependency-aware list scan (gates on disjoint qubits commute): a pass greedily
// absorbs, in program order, every gate whose operands fit in the tile and whose predecessors on
// those qubits are already scheduled; a gate that does not fit blocks its qubits for the rest of the
// scan.  The tile of a pass is the better of that first-fit tile and the best window of consecutive
// qubits (which can take the entangler linking it to the previous window together with the next layer's
// rotations).  Consecutive RX / RZ on one qubit are fused into one 2x2 gate.  For the ring ansatz of
// /root/reference/ROADMAP.md:126-128 on 20 qubits with 12 tile bits this finds the 2-passes-per-layer
// schedule (qubits 0..11 | {0,1,2} + 11..19).  Plain C++ (no device code); it lives in a .cu file
// only so that the one-command build picks it up.
#include <algorithm>
#include <vector>
#include "common.cuh"
#include "../../include/qfedx_b200.h"

namespace {

struct Gate {
    int kind, q0, q1, p0, p1;  // kinds: see QFX_GATE_* in the header; p0/p1 parameter slots (or -1)
};

}  // namespace

// gates_in: n_gates x 4 ints (kind, q0, q1, param).  Outputs (host arrays, caller-allocated):
//   pass_tile_qubits : max_passes x tile_bits   qubit at each tile bit position (ascending)
//   pass_op_begin    : max_passes + 1           offsets into ops_out
//   ops_out          : max_ops x 5 ints         (kind, local bit 0, local bit 1, param 0, param 1)
// Returns the number of passes, or a negative error code.
extern "C" int qfx_circuit_compile(int n_qubits, int n_gates, const int* gates_in, int tile_bits, int low_bits,
                                   int max_passes, int max_ops, int* pass_tile_qubits, int* pass_op_begin,
                                   int* ops_out) {
    QFX_ARG_CHECK(n_qubits >= 1 && n_qubits <= 40 && n_gates >= 0, QFX_ERR_BAD_ARG, "bad circuit size");
    if (tile_bits > n_qubits) tile_bits = n_qubits;
    if (low_bits > tile_bits) low_bits = tile_bits;
    QFX_ARG_CHECK(tile_bits >= 2 || n_qubits < 2, QFX_ERR_BAD_ARG, "tile must hold at least two qubits");

    // 1. fuse RX(q) immediately followed (on that qubit) by RZ(q) into U2
    std::vector<Gate> gates;
    gates.reserve(n_gates);
    std::vector<int> last_on_qubit(n_qubits, -1);
    for (int i = 0; i < n_gates; ++i) {
        Gate g{gates_in[4 * i], gates_in[4 * i + 1], gates_in[4 * i + 2], gates_in[4 * i + 3], -1};
        QFX_ARG_CHECK(g.q0 >= 0 && g.q0 < n_qubits, QFX_ERR_BAD_ARG, "gate %d: qubit %d out of range", i, g.q0);
        const bool two = g.kind == QFX_GATE_CNOT || g.kind == QFX_GATE_CZ;
        if (two) QFX_ARG_CHECK(g.q1 >= 0 && g.q1 < n_qubits && g.q1 != g.q0, QFX_ERR_BAD_ARG, "gate %d: bad second qubit", i);
        if (g.kind == QFX_GATE_RZ && last_on_qubit[g.q0] >= 0 && gates[last_on_qubit[g.q0]].kind == QFX_GATE_RX &&
            gates[last_on_qubit[g.q0]].p1 < 0) {
            Gate& prev = gates[last_on_qubit[g.q0]];
            prev.kind = QFX_GATE_RXRZ;
            prev.p1 = g.p0;
            continue;
        }
        gates.push_back(g);
        last_on_qubit[g.q0] = (int)gates.size() - 1;
        if (two) last_on_qubit[g.q1] = (int)gates.size() - 1;
    }

    // 2. list scheduling into passes
    const int G = (int)gates.size();
    std::vector<char> done(G, 0);
    int n_done = 0, n_pass = 0, n_ops = 0;
    pass_op_begin[0] = 0;
    while (n_done < G) {
        QFX_ARG_CHECK(n_pass < max_passes, QFX_ERR_TOO_LARGE, "more than %d passes", max_passes);
        std::vector<char> in_tile, blocked;
        std::vector<int> chosen;
        int used = 0;
        // One scan over the unscheduled gates in program order.  grow = true: the tile starts with `tile` and
        // absorbs the operands of every gate that still fits (first fit).  grow = false: the tile is fixed and
        // only gates that lie inside it are taken.  A gate that is not taken blocks its qubits for the rest of
        // the scan, which keeps the per-qubit order.
        auto scan = [&](std::vector<char>& tile, int& n_used, bool grow, std::vector<int>& out) {
            blocked.assign(n_qubits, 0);
            out.clear();
            for (int i = 0; i < G; ++i) {
                if (done[i]) continue;
                const Gate& g = gates[i];
                const bool two = g.kind == QFX_GATE_CNOT || g.kind == QFX_GATE_CZ;
                const int need = (tile[g.q0] ? 0 : 1) + ((two && !tile[g.q1]) ? 1 : 0);
                const bool free_qubits = !blocked[g.q0] && !(two && blocked[g.q1]);
                if (free_qubits && (grow ? n_used + need <= tile_bits : need == 0)) {
                    if (!tile[g.q0]) { tile[g.q0] = 1; ++n_used; }
                    if (two && !tile[g.q1]) { tile[g.q1] = 1; ++n_used; }
                    out.push_back(i);
                } else {
                    blocked[g.q0] = 1;
                    if (two) blocked[g.q1] = 1;
                }
            }
        };
        // reserve the low qubits for coalescing; if nothing fits next to them, give some of them up
        for (int lb = low_bits; lb >= 0 && chosen.empty(); --lb) {
            in_tile.assign(n_qubits, 0);
            used = 0;
            for (int q = 0; q < lb; ++q) { in_tile[q] = 1; ++used; }
            scan(in_tile, used, true, chosen);
            // First fit fills the tile with whatever comes first in program order (a whole layer of rotations)
            // and then cannot take the entangler that links it to the previous tile.  Also try every window of
            // consecutive qubits next to the reserved low ones and keep the tile that schedules the most gates.
            const int width = tile_bits - lb;
            if (width >= 1 && n_qubits - lb > width) {
                std::vector<char> cand;
                std::vector<int> cand_chosen;
                for (int start = lb; start < n_qubits; ++start) {
                    cand.assign(n_qubits, 0);
                    int cu = 0;
                    for (int q = 0; q < lb; ++q) { cand[q] = 1; ++cu; }
                    for (int k = 0, q = start; k < n_qubits && cu < tile_bits; ++k, q = (q + 1 == n_qubits ? lb : q + 1))
                        if (!cand[q]) { cand[q] = 1; ++cu; }
                    scan(cand, cu, false, cand_chosen);
                    if (cand_chosen.size() > chosen.size()) {
                        chosen = cand_chosen;
                        in_tile = cand;
                        used = cu;
                    }
                }
            }
        }
        QFX_ARG_CHECK(!chosen.empty(), QFX_ERR_BAD_ARG, "scheduler made no progress (tile too small?)");
        // pad the tile with the lowest unused qubits so every pass moves full-size tiles
        for (int q = 0; q < n_qubits && used < tile_bits; ++q)
            if (!in_tile[q]) { in_tile[q] = 1; ++used; }
        std::vector<int> local_of(n_qubits, -1);
        int pos = 0;
        for (int q = 0; q < n_qubits; ++q)
            if (in_tile[q]) { pass_tile_qubits[n_pass * tile_bits + pos] = q; local_of[q] = pos++; }
        for (int i : chosen) {
            QFX_ARG_CHECK(n_ops < max_ops, QFX_ERR_TOO_LARGE, "more than %d ops", max_ops);
            const Gate& g = gates[i];
            const bool two = g.kind == QFX_GATE_CNOT || g.kind == QFX_GATE_CZ;
            int* o = ops_out + 5 * n_ops++;
            o[0] = g.kind; o[1] = local_of[g.q0]; o[2] = two ? local_of[g.q1] : -1; o[3] = g.p0; o[4] = g.p1;
            done[i] = 1;
            ++n_done;
        }
        pass_op_begin[++n_pass] = n_ops;
    }
    return n_pass;
}
