// lbchem_v2_host.cu — packs a lbchem_mech_desc into the V2Mech blob + Jacobian gather program.
#include <cuda_runtime.h>
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include "../../include/lbchem.h"
#include "lbchem_v2.cuh"

namespace {

struct Blob {
    std::vector<unsigned char> b;
    template <class T>
    int add(const std::vector<T>& v) {
        while (b.size() % 8) b.push_back(0);
        const int off = (int)b.size();
        const unsigned char* p = reinterpret_cast<const unsigned char*>(v.data());
        b.insert(b.end(), p, p + v.size() * sizeof(T));
        while (b.size() % 8) b.push_back(0);
        return off;
    }
};

std::vector<double> soa2(const double* a, int nSp, const int* idx, const double* div, int nq) {
    std::vector<double> v((size_t)nq * nSp);
    for (int i = 0; i < nSp; ++i)
        for (int q = 0; q < nq; ++q) v[(size_t)q * nSp + i] = a[7 * i + idx[q]] / div[q];
    return v;
}

bool small_int(double x, int lo, int hi) { return x == std::floor(x) && x >= lo && x <= hi; }

}  // namespace

// returns 0 = built, 1 = mechanism outside the fast path's class (caller keeps the general kernel), <0 = error

// position of each reaction in the packed tables: classes exp / no exp / irreversible / third body / fall-off
static void reaction_order(const lbchem_mech_desc* m, std::vector<int>& order, std::vector<int>& posOf) {
    const int nR = m->n_reactions;
    order.resize(nR); posOf.resize(nR);
    auto cls = [&](int r) {
        if (m->rtype[r] >= LBCHEM_R_FALLOFF_LINDEMANN) return 4;
        if (m->rtype[r] == LBCHEM_R_THIRD_BODY) return 3;
        if (!m->reversible[r]) return 2;
        const bool hasExp = std::fabs(m->arr[3 * r + 1]) > 1e-300 || std::fabs(m->arr[3 * r + 2]) > 1e-300;
        return hasExp ? 0 : 1;
    };
    for (int r = 0; r < nR; ++r) order[r] = r;
    std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return cls(a) < cls(b); });
    for (int p = 0; p < nR; ++p) posOf[order[p]] = p;
}

// DAC program (chemistryReductionMethods::DAC::reduceMechanism; reference call site LoadBalancedTDACChemistryModel.C:388).
// One int array, offsets in ints from its start:
//   [0] nPairs  [1] pairPtr[nSp+1]  [2] thrTerm[65]  [3] term  [4] pcPtr[nSp+1]  [5] pc  [6] thrPair[65]  [7] nInit
//   [8..15] initial set  [16],[17] tolerance (lo, hi words)
// pairPtr[A]..pairPtr[A+1]: the (A,B) pairs of species A in first-appearance order of B (the rABPos order);
// term: the contributions net_nu_A(reaction) * omega[pos] to the pair numerators, pair by pair, reactions ascending
// (bits 0-15 pos, 16-23 signed net nu, bit 24 closes a pair, bits 25-30 B); thread t of a cell owns the pairs
// [thrPair[t], thrPair[t+1]) = terms [thrTerm[t], thrTerm[t+1]); pc: per species the net_nu_A * omega[pos] of every
// reaction holding A (P_A / C_A).
int lb_dac_program(const lbchem_mech_desc* m, double tol, int nInit, const int32_t* init, std::vector<int>& prog, int* nPairsOut) {
    const int nSp = m->n_species, nR = m->n_reactions;
    if (nSp > 64 || nInit < 1 || nInit > 8 || !(tol > 0.0)) return 1;
    std::vector<int> order, posOf;
    reaction_order(m, order, posOf);
    struct Contribution { int pos, coef; };
    std::vector<std::vector<int>> nbOf(nSp);                                   // B list per A
    std::vector<std::map<int, std::vector<Contribution>>> termsOf(nSp);        // per A: B -> contributions
    std::vector<std::vector<Contribution>> pcOf(nSp);
    for (int r = 0; r < nR; ++r) {
        std::vector<int> ids, net;                                             // species of the reaction, lhs then rhs, each once
        auto add = [&](int s, int nu) {
            for (size_t q = 0; q < ids.size(); ++q) if (ids[q] == s) { net[q] += nu; return; }
            ids.push_back(s); net.push_back(nu);
        };
        for (int k = m->lhs_ptr[r]; k < m->lhs_ptr[r + 1]; ++k) {
            if (m->lhs_nu[k] != std::floor(m->lhs_nu[k]) || m->lhs_nu[k] > 100) return 1;
            add(m->lhs_species[k], -(int)m->lhs_nu[k]);
        }
        for (int k = m->rhs_ptr[r]; k < m->rhs_ptr[r + 1]; ++k) {
            if (m->rhs_nu[k] != std::floor(m->rhs_nu[k]) || m->rhs_nu[k] > 100) return 1;
            add(m->rhs_species[k], (int)m->rhs_nu[k]);
        }
        for (size_t a = 0; a < ids.size(); ++a) {
            const int A = ids[a];
            if (net[a] < -127 || net[a] > 127) return 1;
            for (size_t b = 0; b < ids.size(); ++b) {
                if (b == a) continue;
                const int Bsp = ids[b];
                if (!termsOf[A].count(Bsp)) nbOf[A].push_back(Bsp);
                termsOf[A][Bsp].push_back({posOf[r], net[a]});
            }
            pcOf[A].push_back({posOf[r], net[a]});
        }
    }
    std::vector<int> pairPtr(nSp + 1, 0), pcPtr(nSp + 1, 0), pairTerms;
    std::vector<unsigned> term, pc;
    for (int A = 0; A < nSp; ++A) {
        for (int Bsp : nbOf[A]) {
            const auto& L = termsOf[A][Bsp];
            for (size_t q = 0; q < L.size(); ++q)
                term.push_back((unsigned)L[q].pos | ((unsigned)(uint8_t)(int8_t)L[q].coef << 16) | ((q + 1 == L.size()) ? 1u << 24 : 0u) |
                               ((unsigned)Bsp << 25));
            pairTerms.push_back((int)L.size());
        }
        for (const auto& c : pcOf[A]) pc.push_back((unsigned)c.pos | ((unsigned)(uint8_t)(int8_t)c.coef << 16));
        pairPtr[A + 1] = (int)pairTerms.size(); pcPtr[A + 1] = (int)pc.size();
    }
    // deal the pairs to the 64 threads of a cell in contiguous runs balanced by term count (a pair is never split)
    const int nPairs = (int)pairTerms.size();
    std::vector<int> thrPair(V2_TPC + 1, nPairs), thrTerm(V2_TPC + 1, (int)term.size());
    {
        const double total = (double)term.size();
        double acc = 0; int th = 0, tsum = 0;
        thrPair[0] = 0; thrTerm[0] = 0;
        for (int e = 0; e < nPairs; ++e) {
            while (th + 1 < V2_TPC && acc >= total * (th + 1) / V2_TPC) { ++th; thrPair[th] = e; thrTerm[th] = tsum; }
            acc += pairTerms[e]; tsum += pairTerms[e];
        }
        while (th + 1 <= V2_TPC) { ++th; thrPair[th] = nPairs; thrTerm[th] = tsum; }
    }
    prog.assign(18, 0);
    prog[0] = nPairs;
    auto put = [&](int slot, const void* src, size_t bytes) {
        prog[slot] = (int)prog.size();
        const int* s = static_cast<const int*>(src);
        prog.insert(prog.end(), s, s + bytes / 4);
    };
    if (term.empty()) term.push_back(0);
    if (pc.empty()) pc.push_back(0);
    for (int q = 0; q < 8; ++q) { term.push_back(0); pc.push_back(0); }      // the kernel reads eight words ahead
    put(1, pairPtr.data(), pairPtr.size() * 4); put(2, thrTerm.data(), thrTerm.size() * 4);
    put(3, term.data(), term.size() * 4); put(4, pcPtr.data(), pcPtr.size() * 4); put(5, pc.data(), pc.size() * 4);
    put(6, thrPair.data(), thrPair.size() * 4);
    prog[7] = nInit;
    for (int i = 0; i < nInit; ++i) { if (init[i] < 0 || init[i] >= nSp) return 1; prog[8 + i] = init[i]; }
    std::memcpy(&prog[16], &tol, 8);
    if (nPairsOut) *nPairsOut = pairPtr[nSp];
    return 0;
}

int lb_v2_build(const lbchem_mech_desc* m, V2Mech* out, std::vector<void*>& allocs, std::string& err, std::vector<int>* posOfOut) {
    const int nSp = m->n_species, nR = m->n_reactions, nEff = m->n_eff;
    const int N = nSp + 2;
    if (N > 64 || nSp > 254 || nR > 32000 || nEff > 31 || nR < 1) return 1;
    V2Mech M{};
    M.nSp = nSp; M.nR = nR; M.nEqn = N; M.nEff = nEff;

    // ---- reaction order: homogeneous rounds of 64 (exp / no exp / irreversible / third body / fall-off)
    std::vector<int> order, posOf;
    reaction_order(m, order, posOf);
    if (posOfOut) *posOfOut = posOf;

    std::vector<double> tcTab;
    std::vector<double> A(nR), beta(nR), Ta(nR), fall;
    std::vector<uint32_t> info(nR);
    std::vector<unsigned long long> lpk(nR), rpk(nR);
    int nFall = 0;
    for (int p = 0; p < nR; ++p) {
        const int r = order[p];
        A[p] = m->arr[3 * r]; beta[p] = m->arr[3 * r + 1]; Ta[p] = m->arr[3 * r + 2];
        const int nl = m->lhs_ptr[r + 1] - m->lhs_ptr[r], nr = m->rhs_ptr[r + 1] - m->rhs_ptr[r];
        if (nl < 1 || nr < 1 || nl > 4 || nr > 4) return 1;
        double dn = 0; int sumL = 0, sumR = 0;
        unsigned long long lp = 0, rp = 0;
        for (int s = 0; s < nl; ++s) {
            const int k = m->lhs_ptr[r] + s;
            if (!small_int(m->lhs_nu[k], 1, 7) || !small_int(m->lhs_exp[k], 1, 3)) return 1;
            lp |= (unsigned long long)((unsigned)m->lhs_species[k] | ((unsigned)m->lhs_nu[k] << 10) | ((unsigned)m->lhs_exp[k] << 13)) << (16 * s);
            dn -= m->lhs_nu[k]; sumL += (int)m->lhs_exp[k];
        }
        for (int s = 0; s < nr; ++s) {
            const int k = m->rhs_ptr[r] + s;
            if (!small_int(m->rhs_nu[k], 1, 7) || !small_int(m->rhs_exp[k], 1, 3)) return 1;
            rp |= (unsigned long long)((unsigned)m->rhs_species[k] | ((unsigned)m->rhs_nu[k] << 10) | ((unsigned)m->rhs_exp[k] << 13)) << (16 * s);
            dn += m->rhs_nu[k]; sumR += (int)m->rhs_exp[k];
        }
        if (sumL > 4 || sumR > 4 || !small_int(dn, -3, 3)) return 1;
        lpk[p] = lp; rpk[p] = rp;
        const double rtc = m->Tcommon[m->lhs_species[m->lhs_ptr[r]]];
        int tci = (int)(std::find(tcTab.begin(), tcTab.end(), rtc) - tcTab.begin());
        if (tci == (int)tcTab.size()) tcTab.push_back(rtc);
        if (tci > 7) return 1;
        const int rtype = m->rtype[r];
        int fi = 0, eff = 0;
        if (rtype != LBCHEM_R_ARRHENIUS) {
            eff = m->eff_index[r];
            if (eff < 0 || eff >= nEff) { err = "eff_index out of range"; return LBCHEM_E_INVALID; }
        }
        if (rtype >= LBCHEM_R_FALLOFF_LINDEMANN) {
            if (nFall > 255) return 1;
            fi = nFall++;
            for (int q = 0; q < 3; ++q) fall.push_back(m->arr0[3 * r + q]);
            for (int q = 0; q < 4; ++q) fall.push_back(m->troe[4 * r + q]);
            fall.push_back(0.0);
        }
        const bool hasExp = std::fabs(beta[p]) > 1e-300 || std::fabs(Ta[p]) > 1e-300;
        info[p] = (unsigned)rtype | ((unsigned)(m->reversible[r] ? 1 : 0) << 2) | ((unsigned)nl << 3) | ((unsigned)nr << 6) |
                  ((unsigned)eff << 9) | ((unsigned)tci << 14) | ((unsigned)((int)dn + 3) << 17) | ((unsigned)fi << 20) |
                  ((unsigned)(hasExp ? 1 : 0) << 28);
    }
    M.nFall = nFall; M.nTc = (int)tcTab.size();
    while (tcTab.size() < 8) tcTab.push_back(0.0);

    Blob B;
    M.oW = B.add(std::vector<double>(m->W, m->W + nSp));
    M.oTc = B.add(std::vector<double>(m->Tcommon, m->Tcommon + nSp));
    M.oA = B.add(A); M.oInfo = B.add(info);
    M.oLpk = B.add(lpk); M.oRpk = B.add(rpk);
    if (fall.empty()) fall.push_back(0.0);
    M.oTcTab = B.add(tcTab);

    // sparse third-body efficiencies: M_e = sum_i c_i + sum_{eff != 1} (eff - 1) c_i
    {
        std::vector<int> effPtr(nEff + 1, 0), effSp; std::vector<double> effVal;
        for (int e = 0; e < nEff; ++e) {
            for (int i = 0; i < nSp; ++i) {
                const double v = m->eff[(size_t)e * nSp + i];
                if (v != 1.0) { effSp.push_back(i); effVal.push_back(v - 1.0); }
            }
            effPtr[e + 1] = (int)effSp.size();
        }
        if (effSp.empty()) { effSp.push_back(0); effVal.push_back(0.0); }
        M.oEffPtr = B.add(effPtr); M.oEffSp = B.add(effSp); M.oEffVal = B.add(effVal);
    }

    // omega gather lists: species-major, ORIGINAL reaction order, lhs before rhs (the reference's accumulation order);
    // 16-bit entries (packed position | signed nu << 12), every species' list padded to a multiple of four entries
    // (nu = 0) so that the kernel reads four entries with one 64-bit load
    {
        if (nR > 4095) return 1;
        std::vector<int> spPtr(nSp + 1, 0); std::vector<uint16_t> spList;
        std::vector<std::vector<uint16_t>> L(nSp);
        for (int r = 0; r < nR; ++r) {
            for (int k = m->lhs_ptr[r]; k < m->lhs_ptr[r + 1]; ++k)
                L[m->lhs_species[k]].push_back((uint16_t)((unsigned)posOf[r] | (((unsigned)(-(int)m->lhs_nu[k]) & 15u) << 12)));
            for (int k = m->rhs_ptr[r]; k < m->rhs_ptr[r + 1]; ++k)
                L[m->rhs_species[k]].push_back((uint16_t)((unsigned)posOf[r] | (((unsigned)((int)m->rhs_nu[k]) & 15u) << 12)));
        }
        for (int i = 0; i < nSp; ++i) {
            while (L[i].size() % 4) L[i].push_back(0);
            spList.insert(spList.end(), L[i].begin(), L[i].end());
            spPtr[i + 1] = (int)spList.size();
        }
        if (spList.empty()) spList.assign(4, 0);
        M.oSpPtr = B.add(spPtr); M.oSpList = B.add(spList);
        // Balanced gather: the species sums are dealt to ALL threads of a group. The threads past the species rows
        // (T, p, padding lanes: 11 of 64 for GRI) take parts of the longest lists -- H has 148 terms against a mean of
        // 25 -- so that no lane walks more than ~40. gaTab: [0, 64) first entry, [64, 128) end entry of thread t;
        // [128, 192) owners: first helper slot | helper count << 8 (helper thread nSp + s writes slot s, at most 16 slots); [191] slots in use
        {
            const int tpc = N <= 8 ? 8 : (N <= 16 ? 16 : (N <= 32 ? 32 : 64));
            const int nHelp = std::min(16, std::max(0, tpc - nSp));
            std::vector<int> parts(nSp, 1);
            auto quads = [&](int i) { return (spPtr[i + 1] - spPtr[i]) / 4; };
            auto worst = [&](int i) { return (quads(i) + parts[i] - 1) / parts[i]; };
            for (int h = 0; h < nHelp; ++h) {
                int best = 0;
                for (int i = 1; i < nSp; ++i) if (worst(i) > worst(best)) best = i;
                if (worst(best) <= 6) break;                       // lists of <= 24 terms stay whole (H2/O2, 1-step CH4: no helpers, no barrier)
                parts[best]++;
            }
            std::vector<int> ga(3 * 64, 0);
            int slot = 0;
            for (int i = 0; i < nSp; ++i) {
                const int q = quads(i), n = parts[i], per = (q + n - 1) / n;
                ga[i] = spPtr[i]; ga[64 + i] = spPtr[i] + 4 * std::min(q, per);
                ga[128 + i] = slot | ((n - 1) << 8);
                for (int k = 1; k < n; ++k, ++slot) {
                    const int t = nSp + slot;
                    ga[t] = spPtr[i] + 4 * std::min(q, k * per); ga[64 + t] = spPtr[i] + 4 * std::min(q, (k + 1) * per);
                }
            }
            ga[191] = slot;                                        // helper slots in use (thread 63 never owns a species: nSp <= 62)
            M.oGaTab = B.add(ga);
        }
    }

    // Jacobian: (1) "D terms" d(omega_r)/d(c_col) / nu = k * prod of the other concentrations, one per (reaction, side,
    // species j of that side) -- each is shared by every row the reaction touches; (2) entries (row, col) -> sums of
    // coef * D[term] in the reference's accumulation order (reactions ascending, forward side before reverse side).
    struct Term { int didx; int coef; };
    std::vector<unsigned long long> dterm;                 // D-term program: kidx | s1 << 16 | s2 << 24 | s3 << 32
    std::map<std::pair<int, int>, std::vector<Term>> E;
    for (int r = 0; r < nR; ++r) {
        for (int side = 0; side < 2; ++side) {
            const int32_t* ptr = side ? m->rhs_ptr : m->lhs_ptr;
            const int32_t* spc = side ? m->rhs_species : m->lhs_species;
            const double* ex = side ? m->rhs_exp : m->lhs_exp;
            if (side == 1 && !m->reversible[r]) continue;      // kr == 0: contributes exact zeros
            const int b = ptr[r], e = ptr[r + 1];
            for (int j = b; j < e; ++j) {
                int sp3[3]; int nf = 0;
                for (int i = b; i < e; ++i) {
                    const int reps = (int)ex[i] - (i == j ? 1 : 0);
                    for (int q = 0; q < reps; ++q) { if (nf >= 3) return 1; sp3[nf++] = spc[i]; }
                }
                while (nf < 3) sp3[nf++] = nSp;                // CC[nSp] == 1.0
                Term t; t.didx = (int)dterm.size();
                dterm.push_back((unsigned long long)(2 * posOf[r] + side) | ((unsigned long long)sp3[0] << 16) |
                                ((unsigned long long)sp3[1] << 24) | ((unsigned long long)sp3[2] << 32));
                const int ej = (int)ex[j];
                const int col = spc[j];
                // d(omega)/dc_col = +kf-term for side 0, -kr-term for side 1; species on the lhs lose, rhs gain
                for (int k = m->lhs_ptr[r]; k < m->lhs_ptr[r + 1]; ++k) {
                    Term u = t; u.coef = (side == 0 ? -1 : 1) * (int)m->lhs_nu[k] * ej;
                    E[{m->lhs_species[k], col}].push_back(u);
                }
                for (int k = m->rhs_ptr[r]; k < m->rhs_ptr[r + 1]; ++k) {
                    Term u = t; u.coef = (side == 0 ? 1 : -1) * (int)m->rhs_nu[k] * ej;
                    E[{m->rhs_species[k], col}].push_back(u);
                }
            }
        }
    }
    if (dterm.size() > 4095) return 1;
    M.nD = (int)dterm.size();
    // irreversible reactions still own the (row, col) entries of their rhs columns in the oracle (zeros); they
    // only matter structurally, and a structural zero equals an absent entry for the LU build.
    const int nEnt = (int)E.size();
    M.nEnt = nEnt;
    std::vector<int> rowPtr(N + 1, 0); std::vector<unsigned char> entCol;
    std::vector<int> entTerms;
    {
        int row = 0;
        for (auto& kv : E) {
            while (row < kv.first.first) rowPtr[++row] = (int)entCol.size();
            entCol.push_back((unsigned char)kv.first.second);
            entTerms.push_back((int)kv.second.size());
        }
        while (row < N) rowPtr[++row] = (int)entCol.size();
    }
    // cold tail of the blob (the v3 kernel leaves it in global memory, L2-resident): everything that is read as a batch of
    // independent coalesced loads once per call -- the enthalpy / entropy / cp polynomials (prepare, finish_derivatives),
    // the Arrhenius beta and Ta (prefetched one round ahead by the rate sweep), the fall-off parameters (prefetched before
    // the first pass) -- and the matrix sparsity lists. What stays in shared memory is what every reaction of every sweep
    // reads with data-dependent addresses.
    M.hotBytes = (int)B.b.size();
    {
        const int cpI[5] = {0, 1, 2, 3, 4}; const double cpD[5] = {1, 1, 1, 1, 1};
        const int haI[6] = {4, 3, 2, 1, 0, 5}; const double haD[6] = {5, 4, 3, 2, 1, 1};
        const int sI[6] = {4, 3, 2, 1, 0, 6};  const double sD[6] = {4, 3, 2, 1, 1, 1};
        M.oHaLo = B.add(soa2(m->janaf_low, nSp, haI, haD, 6)); M.oHaHi = B.add(soa2(m->janaf_high, nSp, haI, haD, 6));
        M.oCpLo = B.add(soa2(m->janaf_low, nSp, cpI, cpD, 5)); M.oCpHi = B.add(soa2(m->janaf_high, nSp, cpI, cpD, 5));
        M.oSLo = B.add(soa2(m->janaf_low, nSp, sI, sD, 6));    M.oSHi = B.add(soa2(m->janaf_high, nSp, sI, sD, 6));
        M.oBeta = B.add(beta); M.oTa = B.add(Ta); M.oFall = B.add(fall);
    }
    M.oRowPtr = B.add(rowPtr);
    if (entCol.empty()) entCol.push_back(0);
    M.oEntCol = B.add(entCol);

    // deal entries to `tpc` threads in contiguous runs balanced by term count; program word k of thread t at [k*tpc + t]
    struct Prog { std::vector<unsigned> jterm; std::vector<int> thrEnt, thrN; int maxT; };
    auto deal = [&](int tpc) {
        Prog P;
        std::vector<int> thrEnt(tpc + 1, nEnt), thrN(tpc, 0);
        {
            double total = 0; for (int c : entTerms) total += c + 1.0;
            double acc = 0; int t = 0; thrEnt[0] = 0;
            for (int e = 0; e < nEnt; ++e) {
                while (t + 1 < tpc && acc >= total * (t + 1) / tpc) thrEnt[++t] = e;
                acc += entTerms[e] + 1.0;
            }
            while (t + 1 <= tpc) thrEnt[++t] = nEnt;
        }
        std::vector<std::vector<unsigned>> perThr(tpc);
        {
            int e = 0;
            for (auto& kv : E) {
                int t = 0; while (!(e >= thrEnt[t] && e < thrEnt[t + 1])) ++t;
                for (size_t q = 0; q < kv.second.size(); ++q) {
                    const Term& u = kv.second[q];
                    unsigned w = (unsigned)u.didx | ((unsigned)(uint8_t)(int8_t)u.coef << 12);
                    if (q + 1 == kv.second.size()) w |= 1u << 20;
                    perThr[t].push_back(w);
                }
                ++e;
            }
        }
        int maxT = 1;
        for (int t = 0; t < tpc; ++t) { thrN[t] = (int)perThr[t].size(); maxT = std::max(maxT, thrN[t]); }
        std::vector<unsigned> jterm((size_t)(maxT + 16) * tpc, 0u);   // + 16 rows: the kernel reads eight words ahead of a batch of eight
        for (int t = 0; t < tpc; ++t)
            for (int k = 0; k < thrN[t]; ++k) jterm[(size_t)k * tpc + t] = perThr[t][k];


        P.jterm = jterm; P.thrEnt = thrEnt; P.thrN = thrN; P.maxT = maxT;
        return P;
    };
    // group width: the smallest of 8 / 16 / 32 / 64 threads that gives every equation its own thread
    M.tpc = N <= 8 ? 8 : (N <= 16 ? 16 : (N <= 32 ? 32 : 64));
    if (const char* e = getenv("LBCHEM_V3_TPC")) { const int q = atoi(e); if ((q == 8 || q == 16 || q == 32 || q == 64) && q >= N) M.tpc = q; }
    Prog P64 = deal(M.tpc);
    M.jMaxTerms = P64.maxT;
    M.blobBytes = (int)B.b.size();
    auto up = [&](const void* src, size_t bytes, const void** dst) -> bool {
        void* d = nullptr;
        if (cudaMalloc(&d, bytes ? bytes : 8) != cudaSuccess) return false;
        allocs.push_back(d);
        if (bytes && cudaMemcpy(d, src, bytes, cudaMemcpyHostToDevice) != cudaSuccess) return false;
        *dst = d;
        return true;
    };
    const void *dBlob, *dTerm, *dEnt, *dN, *dD;
    for (int q = 0; q < 8 * 64; ++q) dterm.push_back(0ull);          // read-ahead padding
    if (!up(B.b.data(), B.b.size(), &dBlob) || !up(P64.jterm.data(), P64.jterm.size() * 4, &dTerm) || !up(dterm.data(), dterm.size() * 8, &dD) ||
        !up(P64.thrEnt.data(), P64.thrEnt.size() * 4, &dEnt) || !up(P64.thrN.data(), P64.thrN.size() * 4, &dN)) {
        err = "cudaMalloc/cudaMemcpy failed while uploading the packed mechanism";
        return LBCHEM_E_CUDA;
    }
    M.blob = static_cast<const unsigned char*>(dBlob);
    M.jterm = static_cast<const unsigned*>(dTerm);
    M.dterm = static_cast<const unsigned long long*>(dD);
    M.jthrEnt = static_cast<const int*>(dEnt);
    M.jthrN = static_cast<const int*>(dN);
    *out = M;
    return 0;
}

// per-cell workspace (doubles) of the fast path; must match carve_v2 in lbchem_kernels_v2.cu
int lb_v2_workspace(V2Mech& M, int solver, double relTol, double absTol) {
    const int N = M.nEqn, nSp = M.nSp;
    int tb = 0;
    if (solver == LBCHEM_SOLVER_SEULEX) {
        int kt = (int)(-std::log10(relTol + absTol) * 0.6 + 0.5);
        kt = std::max(1, std::min(10, kt));
        tb = std::min(11, kt + 4);
    }
    M.tbRows = tb;
    const int nBlk = (N + V2_BLK - 1) / V2_BLK;
    int luAux = nBlk * (V2_LSTRIDE + V2_USTRIDE) + 4 * (N + 2);      // LB, UB, pivot-row buffers [2 warps][2 parities]
    if (luAux < 2 * M.nR) luAux = 2 * M.nR;                            // aliased by KF/KR during the Jacobian
    int d = 10 * N + tb * N + (nSp + 1) + 2 * nSp + M.nEff + M.nR + M.nEnt + nSp   // vectors, TB, CC, GL, GH, ME, OM, JS, JT
            + 2 * N + 16 + luAux + 16 + 16;                                         // RD, BSH, XS, luAux, exchange slots, reduction scratch
    d = (d + 1) & ~1;
    M.wsDoubles = d;
    return d;
}
