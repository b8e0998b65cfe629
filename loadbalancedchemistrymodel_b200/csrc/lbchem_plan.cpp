// lbchem_plan.cpp — the cost-balancing plan of LoadBalancedChemistryModel, host side, bit-exact with
// the reference for the same cost vector.
//   getSortedCPUTimesOnProcessor  reference LoadBalancedChemistryModel.C:311-347
//   updateProcessorBalancing      reference LoadBalancedChemistryModel.C:170-308
//   (TDAC variant)                reference LoadBalancedTDACChemistryModel.C:149-287
//   cellsToSend                   reference LoadBalancedChemistryModel.C:143-167
#include <algorithm>
#include <functional>
#include <tuple>
#include <utility>
#include <vector>

#include "../../include/lbchem.h"

namespace {
struct Load {
    double t;                   // (remaining) cost of the rank
    std::pair<int, int> id;     // (procID, nCells): the reference's Pair<label>
};
struct Send { int to; double amount; };
}  // namespace

extern "C" int lbchem_balance_plan(int32_t nranks, const double* cost, const int32_t* cells, double min_fraction,
                                   int32_t tdac, int32_t* send_ptr, int32_t* send_to, double* send_frac,
                                   int32_t* recv_ptr, int32_t* recv_from) {
    if (nranks < 1 || !cost || !send_ptr || !send_to || !send_frac || !recv_ptr || !recv_from) return LBCHEM_E_INVALID;
    const int P = nranks;
    std::vector<Load> S(P);
    for (int r = 0; r < P; ++r) S[r] = Load{cost[r], {r, cells ? cells[r] : 0}};
    // std::greater on pair<scalar, Pair<label>>: descending cost, ties -> larger procID, then larger nCells
    std::stable_sort(S.begin(), S.end(), [](const Load& a, const Load& b) {
        return std::make_pair(a.t, a.id) > std::make_pair(b.t, b.id);
    });
    double avg = 0.0;
    for (const Load& l : S) avg += l.t;       // summed in sorted order, like the reference
    avg /= static_cast<double>(P);

    std::vector<std::vector<Send>> sends(P);
    std::vector<std::vector<int>> recvs(P);
    for (int i = 0; i < P; ++i) {
        const int from = S[i].id.first;
        const double tFrom = S[i].t;          // captured once
        double surplus = tFrom - avg;
        if (tdac && surplus < 0) continue;
        for (int k = P - 1; k > 0; --k) {     // sorted index 0 never receives
            const int to = S[k].id.first;
            const double room = avg - S[k].t; // live value
            if (room <= 0) continue;
            const double left = room - surplus;
            if (std::min(room, surplus) / tFrom < min_fraction) continue;
            if (left > 0) {
                S[k].t += surplus; S[i].t -= surplus;
                recvs[to].push_back(from);
                sends[from].push_back(Send{to, tdac ? surplus : surplus / tFrom});
                surplus = 0;
                break;
            }
            S[k].t += room; S[i].t -= room; surplus -= room;
            recvs[to].push_back(from);
            sends[from].push_back(Send{to, tdac ? room : room / tFrom});
        }
    }
    int a = 0, b = 0;
    for (int r = 0; r < P; ++r) {
        send_ptr[r] = a; recv_ptr[r] = b;
        for (const Send& s : sends[r]) { send_to[a] = s.to; send_frac[a] = s.amount; ++a; }
        for (int f : recvs[r]) recv_from[b++] = f;
    }
    send_ptr[P] = a; recv_ptr[P] = b;
    return LBCHEM_OK;
}

extern "C" int32_t lbchem_cells_to_send(const double* cell_cost, int32_t n, double target, int32_t start) {
    double acc = 0.0;
    for (int32_t i = start; i < n; ++i) {
        if (acc >= target) return i;
        acc += cell_cost[i];
    }
    return n;
}
