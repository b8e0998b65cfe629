// lbchem_v2.cuh — packed mechanism of the fast path (integrate_v2_kernel): GRI-class mechanisms
// (integer exponents 1..3, <= 4 species per side, nEqn <= 64), 64 threads per cell.
//
// The hot coefficient tables live in ONE blob that every CTA stages into shared memory once
// (persistent kernel); only the Jacobian gather program is streamed from global memory (L2).
#pragma once
#include <stdint.h>

#define V2_TPC 64
#define V2_BLK 8           // column block of the triangular solves
#define V2_LSTRIDE 28      // strictly-lower 8x8 triangle
#define V2_USTRIDE 36      // upper 8x8 triangle incl. diagonal

// info word of a reaction
#define V2_RTYPE(w) ((w) & 3u)
#define V2_REV(w) (((w) >> 2) & 1u)
#define V2_NL(w) (((w) >> 3) & 7u)
#define V2_NR(w) (((w) >> 6) & 7u)
#define V2_EFF(w) (((w) >> 9) & 31u)
#define V2_TCI(w) (((w) >> 14) & 7u)
#define V2_DNC(w) (((w) >> 17) & 7u)      // dn + 3
#define V2_FALL(w) (((w) >> 20) & 255u)
#define V2_HASEXP(w) (((w) >> 28) & 1u)
// stoichiometry slot (16 bits): species | nu << 10 | exp << 13
#define V2_SLOT(pk, s) ((unsigned)((pk) >> (16 * (s))) & 0xffffu)
#define V2_SP(sl) ((sl) & 1023u)
#define V2_NU(sl) (((sl) >> 10) & 7u)
#define V2_EX(sl) (((sl) >> 13) & 7u)

struct V2Mech {
    int nSp, nR, nEqn, nEff, nFall, nEnt, nTc;
    int blobBytes;
    int hotBytes;                        // leading part of the blob that the v3 kernel stages in shared memory
    // byte offsets into the blob (all 8-byte aligned)
    int oW, oTc, oCpLo, oCpHi, oHaLo, oHaHi, oSLo, oSHi;
    int oA, oBeta, oTa, oInfo, oLpk, oRpk, oFall, oTcTab;
    int oEffPtr, oEffSp, oEffVal;
    int oSpPtr, oSpList, oGaTab;
    int oRowPtr, oEntCol;
    const unsigned char* blob;           // global copy of the blob
    // Jacobian: D-term program [nD] (k index and the three concentration factors of d omega / d c), then the entry gather
    // program (global, thread-interleaved: term k of thread t at [k*64 + t]; 32-bit words: D index, coefficient, last flag)
    const unsigned long long* dterm; int nD;
    const unsigned* jterm;
    const int* jthrEnt;                  // [65] first entry of thread t
    const int* jthrN;                    // [64] number of terms of thread t
    int jMaxTerms;
    const int* dac;                      // DAC program (lb_dac_program) when reduction is active, else null (v3 kernel)
    int tpc;                             // threads per cell of the compact-code kernel: 8, 16, 32 or 64 (>= nEqn)
    int tbRows;                          // rows of the seulex extrapolation table held per cell
    int wsDoubles;                       // per-cell workspace in doubles
    int rlInMatrix;                      // DAC: the enabled-reaction list lives in the tail of the matrix region (set by lbchem_set_reduction)
};

// D term (64 bits): [0,16) index into KFR (2*pos + side)   [16,24) [24,32) [32,40) species factors (nSp = the constant 1)
#define V2_T_K(w) ((unsigned)((w) & 0xffffull))
#define V2_T_S1(w) ((unsigned)(((w) >> 16) & 0xffull))
#define V2_T_S2(w) ((unsigned)(((w) >> 24) & 0xffull))
#define V2_T_S3(w) ((unsigned)(((w) >> 32) & 0xffull))
// Jacobian gather term (32 bits): [0,12) D-term index   [12,20) signed coefficient (-/+ nu * e_j)   bit 20: last term of its entry
#define V2_J_D(w) ((w) & 0xfffu)
#define V2_J_COEF(w) ((int)(signed char)(((w) >> 12) & 0xffu))
#define V2_J_LAST(w) (((w) >> 20) & 1u)
