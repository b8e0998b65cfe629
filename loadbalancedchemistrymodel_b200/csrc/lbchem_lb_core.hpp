// lbchem_lb_core.hpp -- one flow step of the load-balanced model, independent of where the cells live.
//
// Mirrors LoadBalancedChemistryModel::solve<DeltaTType>
//   reference src/chemistryModel/LoadBalancedChemistryModel/LoadBalancedChemistryModel.C
//     :471-509  firstTime_: everything is solved locally, the per-cell cost is recorded
//     :515-519  re-plan every updateIter + 1 steps (updateProcessorBalancing, :170-347 -> lbchem_balance_plan)
//     :521-570  ship the FIRST cells of the list to each planned receiver (cellsToSend, :143-167), keep the tail
//     :572-611  receive; solve own tail first, then the received cells
//     :613-673  results go back to their owners in the same order; :676-701 unpack, total cost
//   and its LoadBalancedTDAC variant (LoadBalancedTDACChemistryModel.C:623-979: absolute cost targets :232,256,
//   literal 2 % threshold :214, deltaTChem never copied back :968-971),
//   with src/pointToPointBuffer/pointToPointBuffer.C:53-168 replaced by a Transport.
//
// What differs from the reference: a cell is a COLUMN of field-major blocks (Y[i][cell], OpenFOAM's own layout), so a
// contiguous cell range [s, e) is nSp + 5 contiguous row slices: they are sent as they lie (no serialisation, no packing)
// and received straight into a field-major block on the other side; results come back straight into the owner's RR rows.
// Sizes are known to both sides from one all-gather of the per-destination counts (the reference's 8-byte size
// handshake, pointToPointBuffer.C:194-267); the return trip needs none (:53-66).
//
// The step logic is written against two small interfaces so that the SAME code runs on the device (CUDA backend, NCCL or
// caller-supplied transport: lbchem_lb_cuda.cu) and on the host in the CPU tests (plain memory, a stand-in cell solver,
// gloo behind the transport callbacks: lbchem_lb_host.cpp).
#pragma once
#include <algorithm>
#include <condition_variable>
#include <cstdint>
#include <cstring>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/lbchem.h"

namespace lbcore {

struct Transport {
    int rank = 0, size = 1;
    std::string err;
    virtual ~Transport() {}
    // blocking all-gather of `count` doubles per rank; HOST buffers
    virtual int all_gather(const double* send, double* recv, int count) = 0;
    // one grouped point-to-point exchange; ops[].ptr are BACKEND memory. Ops between one pair of ranks are matched in
    // posting order. `stream`: the step's stream (cudaStream_t) -- an asynchronous transport (NCCL) enqueues on it, a
    // blocking one is called with the stream already drained and returns when the data has arrived.
    virtual int exchange(const std::vector<lbchem_p2p_op>& ops, void* stream) = 0;
    virtual bool async() const { return false; }
};

struct Block {   // field-major cells in backend memory (the reference's cell list after updateCellDataList)
    long long n = 0, ldy = 0;
    const double *Y = nullptr, *T = nullptr, *p = nullptr, *rho = nullptr, *deltaT = nullptr;
    double *deltaTChem = nullptr, *RR = nullptr, *Tout = nullptr, *cost = nullptr;
    int32_t* status = nullptr;
};

struct Backend {
    std::string err;
    virtual ~Backend() {}
    virtual int n_species() const = 0;
    virtual void* alloc(size_t bytes) = 0;
    virtual void release(void* p) = 0;
    virtual int copy(void* dst, const void* src, size_t bytes, void* stream) = 0;     // backend memory, stream-ordered
    virtual int sync(void* stream) = 0;
    // solveCellList(own cells [aStart, a.n)) then solveCellList(b) (reference .C:608-611); stream-ordered
    virtual int solve(const Block& a, long long aStart, const Block* b, double deltaTUniform, void* stream) = 0;
    // updateTotalCpuTime (.C:350-362) + min(deltaTChem) (.C:687-693) over n cells, results on the HOST (blocking).
    // Also prepares whatever bounds() needs to answer cellsToSend queries on this same cost vector.
    virtual int reduce(const double* cost, const double* dtc, long long n, double* total, double* dtMin, void* stream) = 0;
    // cellsToSend (.C:143-167) for nT targets in sequence: bounds[0] = 0, bounds[k+1] = first index >= bounds[k] at which
    // the cost summed from bounds[k] reaches targets[k] (else n). HOST in/out, blocking. `cost` is the vector of reduce().
    virtual int bounds(const double* cost, long long n, const double* targets, int nT, long long* bounds, void* stream) = 0;
};

// In-process transport between the ranks of one address space (one host thread per rank): the forward / return
// messages are plain backend copies (device-to-device on one GPU or between peer GPUs). Stands in for NCCL where two
// logical ranks share one GPU (tests), and serves a single-process multi-GPU host.
struct LoopHub {
    int size;
    std::mutex m; std::condition_variable cv; int arrived = 0; long gen = 0;
    std::vector<const double*> ag;
    std::vector<const std::vector<lbchem_p2p_op>*> posted;
    explicit LoopHub(int n) : size(n), ag(n, nullptr), posted(n, nullptr) {}
    void barrier() {
        std::unique_lock<std::mutex> lk(m);
        const long g = gen;
        if (++arrived == size) { arrived = 0; ++gen; cv.notify_all(); }
        else cv.wait(lk, [&] { return gen != g; });
    }
};

struct LoopTransport : Transport {
    std::shared_ptr<LoopHub> hub; Backend* be;
    LoopTransport(std::shared_ptr<LoopHub> h, int r, Backend* b) : hub(std::move(h)), be(b) { rank = r; size = hub->size; }
    int all_gather(const double* send, double* recv, int count) override {
        hub->ag[rank] = send;
        hub->barrier();
        for (int r = 0; r < size; ++r) std::memcpy(recv + (size_t)r * count, hub->ag[r], sizeof(double) * count);
        hub->barrier();
        return LBCHEM_OK;
    }
    int exchange(const std::vector<lbchem_p2p_op>& ops, void* stream) override {
        hub->posted[rank] = &ops;
        hub->barrier();
        std::vector<size_t> cur(size, 0);
        int rc = LBCHEM_OK;
        for (const lbchem_p2p_op& r : ops) {
            if (r.is_send) continue;
            const std::vector<lbchem_p2p_op>& theirs = *hub->posted[r.peer];
            size_t& c = cur[r.peer];
            while (c < theirs.size() && !(theirs[c].is_send && theirs[c].peer == rank)) ++c;
            if (c == theirs.size() || theirs[c].bytes != r.bytes) { err = "loopback exchange: unmatched receive"; rc = LBCHEM_E_INVALID; break; }
            if (rc == LBCHEM_OK && r.bytes) rc = be->copy(r.ptr, theirs[c].ptr, (size_t)r.bytes, stream);
            ++c;
        }
        if (rc == LBCHEM_OK) rc = be->sync(stream);
        hub->barrier();
        return rc;
    }
};

// caller-supplied transport (MPI in the OpenFOAM adapter, gloo in the CPU tests)
struct CallbackTransport : Transport {
    lbchem_transport cb;
    CallbackTransport(const lbchem_transport& c, int r, int n) : cb(c) { rank = r; size = n; }
    int all_gather(const double* send, double* recv, int count) override {
        const int rc = cb.all_gather(cb.ctx, send, recv, count);
        if (rc) err = "transport all_gather callback failed";
        return rc ? LBCHEM_E_INVALID : LBCHEM_OK;
    }
    int exchange(const std::vector<lbchem_p2p_op>& ops, void*) override {
        const int rc = ops.empty() ? 0 : cb.exchange(cb.ctx, (int32_t)ops.size(), ops.data());
        if (rc) err = "transport exchange callback failed";
        return rc ? LBCHEM_E_INVALID : LBCHEM_OK;
    }
};

class Driver {
public:
    Driver(std::unique_ptr<Backend> be, std::unique_ptr<Transport> tp, const lbchem_lb_options& o)
        : be_(std::move(be)), tp_(std::move(tp)), opt_(o), iter_(o.update_iter) {}   // .C:58: plan on the first balanced step
    ~Driver() {
        void* ptrs[] = {lastCost_, rY_, rRR_, rT_, rp_, rrho_, rdT_, rdtc_, rTout_, rcost_, rstatus_, oTout_, ocost_, ostatus_, tdtc_};
        for (void* p : ptrs) if (p) be_->release(p);
    }
    Backend* backend() { return be_.get(); }
    Transport* transport() { return tp_.get(); }
    const std::string& error() const { return err_; }
    const lbchem_lb_info& info() const { return info_; }
    const std::vector<int32_t>& send_to() const { return sendTo_; }
    const std::vector<double>& send_amount() const { return sendAmt_; }
    const std::vector<int32_t>& recv_from() const { return recvFrom_; }
    const std::vector<long long>& bounds() const { return bounds_; }
    const std::vector<long long>& recv_sizes() const { return recvSizes_; }
    void set_options(const lbchem_lb_options& o) { opt_ = o; }
    void reset() { firstTime_ = true; iter_ = opt_.update_iter; lastN_ = -1; }

    // One flow step over this rank's cells. Returns LBCHEM_OK or a negative code; *dtMin = min chemical time step.
    int solve(Block own, double deltaTUniform, double* dtMin, void* stream) {
        const int nSp = be_->n_species();
        const long long n = own.n;
        info_ = lbchem_lb_info{};
        info_.rank = tp_->rank; info_.nranks = tp_->size;
        if (!ensure(oTout_, oCap_[0], n, 8) || !ensure(ocost_, oCap_[1], n, 8) || !ensure(ostatus_, oCap_[2], n, 4) ||
            !ensure(lastCost_, lastCap_, n, 8, true))
            return fail(LBCHEM_E_NOMEM, "out of backend memory");
        if (!own.Tout) own.Tout = static_cast<double*>(oTout_);
        if (!own.cost) own.cost = static_cast<double*>(ocost_);
        if (!own.status) own.status = static_cast<int32_t*>(ostatus_);
        double* callerDtc = own.deltaTChem;
        if (opt_.tdac) {
            // LoadBalancedTDAC never copies cData.deltaTChem() back into deltaTChem_ (.C:968-971): every flow
            // step starts from the caller's field, which stays as it is
            if (!ensure(tdtc_, tdCap_, n, 8)) return fail(LBCHEM_E_NOMEM, "out of backend memory");
            if (n && be_->copy(tdtc_, callerDtc, (size_t)n * 8, stream)) return fail(LBCHEM_E_CUDA, be_->err);
            own.deltaTChem = static_cast<double*>(tdtc_);
        }

        // every rank must take the same branch: the counts all-gather below carries a validity flag
        bool exchange = opt_.balance && tp_->size > 1 && !firstTime_;
        long long start = 0;
        if (exchange) {
            const int P = tp_->size;
            const bool mine_ok = lastN_ == n;
            if (iter_ >= opt_.update_iter) {                      // .C:515-519 (iter_++ >= maxIterUpdate_)
                int rc = plan(n);
                if (rc != LBCHEM_OK) return rc;
                iter_ = 0;
                info_.planned = 1;
            } else ++iter_;
            // ---- which cells go where: cellsToSend on last step's per-cell cost (.C:528-562)
            const int nS = (int)sendTo_.size();
            bounds_.assign(nS + 1, 0);
            if (nS && mine_ok) {
                std::vector<double> targets(nS);
                for (int k = 0; k < nS; ++k) targets[k] = opt_.tdac ? sendAmt_[k] : totalCost_ * sendAmt_[k];   // .C:538 / TDAC .C:781
                if (be_->bounds(static_cast<double*>(lastCost_), n, targets.data(), nS, bounds_.data(), stream))
                    return fail(LBCHEM_E_CUDA, be_->err);
            }
            // ---- size exchange (pointToPointBuffer.C:194-267 sends 8 bytes to each partner): one all-gather of the
            //      per-destination counts, plus "my cost vector matches my cell list"
            std::vector<double> counts(P + 1, 0.0), all((size_t)P * (P + 1));
            for (int k = 0; k < nS; ++k) counts[sendTo_[k]] = (double)(bounds_[k + 1] - bounds_[k]);
            counts[P] = mine_ok ? 1.0 : 0.0;
            if (tp_->all_gather(counts.data(), all.data(), P + 1)) return fail(LBCHEM_E_INVALID, tp_->err);
            bool all_ok = true;
            for (int r = 0; r < P; ++r) all_ok = all_ok && all[(size_t)r * (P + 1) + P] == 1.0;
            if (!all_ok) {
                exchange = false;                                 // a cell list changed size: solve locally, re-plan next step
                iter_ = opt_.update_iter;
            } else {
                recvSizes_.assign(recvFrom_.size(), 0);
                long long mtot = 0;
                for (size_t q = 0; q < recvFrom_.size(); ++q) {
                    recvSizes_[q] = (long long)all[(size_t)recvFrom_[q] * (P + 1) + tp_->rank];
                    mtot += recvSizes_[q];
                }
                start = bounds_[nS];
                int rc = exchange_and_solve(own, deltaTUniform, nSp, start, mtot, stream);
                if (rc != LBCHEM_OK) return rc;
            }
        }
        if (!exchange) {
            if (n && be_->solve(own, 0, nullptr, deltaTUniform, stream)) return fail(LBCHEM_E_CUDA, be_->err);
            firstTime_ = false;
        }
        // ---- unpack + updateTotalCpuTime (.C:676-701, :350-362)
        if (n && be_->copy(lastCost_, own.cost, (size_t)n * 8, stream)) return fail(LBCHEM_E_CUDA, be_->err);
        double total = 0.0, mn = 1e15;
        if (be_->reduce(static_cast<double*>(lastCost_), opt_.tdac ? callerDtc : own.deltaTChem, n, &total, &mn, stream))
            return fail(LBCHEM_E_CUDA, be_->err);
        totalCost_ = total; lastN_ = n;
        info_.total_cost = total; info_.n_cells = n;
        if (dtMin) *dtMin = n ? std::min(mn, 1e15) : 1e15;        // GREAT when the rank holds no cells (.C:455)
        return LBCHEM_OK;
    }

private:
    bool ensure(void*& p, long long& cap, long long n, int elem, bool keep = false) {
        if (n <= cap) return true;
        const long long ncap = n + n / 4 + 64;
        void* q = be_->alloc((size_t)ncap * elem);
        if (!q) return false;
        (void)keep;
        if (p) be_->release(p);                 // (the CUDA backend's release drains the device first)
        p = q; cap = ncap;
        return true;
    }
    int fail(int code, const std::string& m) { err_ = m; return code; }

    // getSortedCPUTimesOnProcessor + updateProcessorBalancing (.C:170-347): every rank derives the whole plan from the
    // all-gathered (total cost, cell count) vector with the bit-exact host function
    int plan(long long n) {
        const int P = tp_->size;
        double mine[2] = {totalCost_, (double)n};
        std::vector<double> all(2 * (size_t)P);
        if (tp_->all_gather(mine, all.data(), 2)) return fail(LBCHEM_E_INVALID, tp_->err);
        std::vector<double> cost(P); std::vector<int32_t> cells(P);
        for (int r = 0; r < P; ++r) { cost[r] = all[2 * r]; cells[r] = (int32_t)all[2 * r + 1]; }
        std::vector<int32_t> sp(P + 1), rp(P + 1), st((size_t)P * P), rf((size_t)P * P);
        std::vector<double> sf((size_t)P * P);
        // LoadBalancedTDAC hard-codes its 2 % threshold (LoadBalancedTDACChemistryModel.C:214)
        const double minFrac = opt_.tdac ? 0.02 : opt_.min_fraction;
        const int rc = lbchem_balance_plan(P, cost.data(), cells.data(), minFrac, opt_.tdac, sp.data(), st.data(), sf.data(), rp.data(), rf.data());
        if (rc != LBCHEM_OK) return fail(rc, "lbchem_balance_plan failed");
        const int me = tp_->rank;
        sendTo_.assign(st.begin() + sp[me], st.begin() + sp[me + 1]);
        sendAmt_.assign(sf.begin() + sp[me], sf.begin() + sp[me + 1]);
        recvFrom_.assign(rf.begin() + rp[me], rf.begin() + rp[me + 1]);
        gathered_ = cost;
        return LBCHEM_OK;
    }

    static void op(std::vector<lbchem_p2p_op>& v, int isSend, int peer, const void* p, long long bytes) {
        lbchem_p2p_op o; o.is_send = isSend; o.peer = peer; o.ptr = const_cast<void*>(p); o.bytes = bytes;
        v.push_back(o);
    }

    int exchange_and_solve(const Block& own, double deltaTUniform, int nSp, long long start, long long mtot, void* stream) {
        const bool perCellDt = own.deltaT != nullptr;      // LTS: a deltaT row travels with the cells (all ranks alike)
        // ---- receive block [nSp][rcap_] (+ per-cell vectors), grown with head-room: no allocation once sizes settle
        if (mtot > rcap_) {
            const long long ncap = mtot + mtot / 4 + 1024;
            be_->sync(stream);
            void** ptrs[] = {&rY_, &rRR_, &rT_, &rp_, &rrho_, &rdT_, &rdtc_, &rTout_, &rcost_, &rstatus_};
            const size_t bytes[] = {(size_t)nSp * 8, (size_t)nSp * 8, 8, 8, 8, 8, 8, 8, 8, 4};
            for (int q = 0; q < 10; ++q) {
                if (*ptrs[q]) be_->release(*ptrs[q]);
                *ptrs[q] = be_->alloc(bytes[q] * (size_t)ncap);
                if (!*ptrs[q]) { rcap_ = 0; return fail(LBCHEM_E_NOMEM, "out of backend memory for the receive block"); }
            }
            rcap_ = ncap;
            info_.reallocated = 1;
        }
        Block rb;
        rb.n = mtot; rb.ldy = rcap_;
        rb.Y = static_cast<double*>(rY_); rb.T = static_cast<double*>(rT_); rb.p = static_cast<double*>(rp_);
        rb.rho = static_cast<double*>(rrho_); rb.deltaT = perCellDt ? static_cast<double*>(rdT_) : nullptr;
        rb.deltaTChem = static_cast<double*>(rdtc_); rb.RR = static_cast<double*>(rRR_); rb.Tout = static_cast<double*>(rTout_);
        rb.cost = static_cast<double*>(rcost_); rb.status = static_cast<int32_t*>(rstatus_);

        // ---- forward trip (.C:528-606): row slices of the first cells, as they lie
        ops_.clear();
        const int nS = (int)sendTo_.size();
        long long sentCells = 0, fwdRows = nSp + 4 + (perCellDt ? 1 : 0);
        for (int k = 0; k < nS; ++k) {
            const long long s = bounds_[k], cnt = bounds_[k + 1] - s;
            if (cnt <= 0) continue;
            const int to = sendTo_[k];
            for (int i = 0; i < nSp; ++i) op(ops_, 1, to, own.Y + (size_t)i * own.ldy + s, cnt * 8);
            op(ops_, 1, to, own.T + s, cnt * 8); op(ops_, 1, to, own.p + s, cnt * 8); op(ops_, 1, to, own.rho + s, cnt * 8);
            if (perCellDt) op(ops_, 1, to, own.deltaT + s, cnt * 8);
            op(ops_, 1, to, own.deltaTChem + s, cnt * 8);
            sentCells += cnt;
        }
        {
            long long off = 0;
            for (size_t q = 0; q < recvFrom_.size(); ++q) {
                const long long m = recvSizes_[q];
                if (m <= 0) continue;
                const int src = recvFrom_[q];
                for (int i = 0; i < nSp; ++i) op(ops_, 0, src, rb.Y + (size_t)i * rb.ldy + off, m * 8);
                op(ops_, 0, src, rb.T + off, m * 8); op(ops_, 0, src, rb.p + off, m * 8); op(ops_, 0, src, rb.rho + off, m * 8);
                if (perCellDt) op(ops_, 0, src, rb.deltaT + off, m * 8);
                op(ops_, 0, src, rb.deltaTChem + off, m * 8);
                off += m;
            }
        }
        // (collective: every rank calls it, also with nothing to ship, so that in-process transports can rendezvous)
        if (!tp_->async() && be_->sync(stream)) return fail(LBCHEM_E_CUDA, be_->err);
        if (tp_->exchange(ops_, stream)) return fail(LBCHEM_E_INVALID, tp_->err);
        // ---- own tail first, then the received cells (.C:608-611): one launch, one work queue
        if ((own.n - start > 0 || mtot > 0) && be_->solve(own, start, mtot > 0 ? &rb : nullptr, deltaTUniform, stream))
            return fail(LBCHEM_E_CUDA, be_->err);
        // ---- return trip (.C:613-673): same order, sizes swapped (pointToPointBuffer.C:53-66), straight into the owner's rows
        ops_.clear();
        {
            long long off = 0;
            for (size_t q = 0; q < recvFrom_.size(); ++q) {
                const long long m = recvSizes_[q];
                if (m <= 0) continue;
                const int src = recvFrom_[q];
                for (int i = 0; i < nSp; ++i) op(ops_, 1, src, rb.RR + (size_t)i * rb.ldy + off, m * 8);
                op(ops_, 1, src, rb.deltaTChem + off, m * 8); op(ops_, 1, src, rb.Tout + off, m * 8);
                op(ops_, 1, src, rb.cost + off, m * 8); op(ops_, 1, src, rb.status + off, m * 4);
                off += m;
            }
        }
        for (int k = 0; k < nS; ++k) {
            const long long s = bounds_[k], cnt = bounds_[k + 1] - s;
            if (cnt <= 0) continue;
            const int to = sendTo_[k];
            for (int i = 0; i < nSp; ++i) op(ops_, 0, to, own.RR + (size_t)i * own.ldy + s, cnt * 8);
            op(ops_, 0, to, own.deltaTChem + s, cnt * 8); op(ops_, 0, to, own.Tout + s, cnt * 8);
            op(ops_, 0, to, own.cost + s, cnt * 8); op(ops_, 0, to, own.status + s, cnt * 4);
        }
        if (!tp_->async() && be_->sync(stream)) return fail(LBCHEM_E_CUDA, be_->err);
        if (tp_->exchange(ops_, stream)) return fail(LBCHEM_E_INVALID, tp_->err);
        info_.exchanged = 1;
        info_.sent_cells = sentCells; info_.recv_cells = mtot;
        info_.sent_bytes = sentCells * 8 * fwdRows + mtot * (8 * (nSp + 3) + 4);        // forward of mine + return of theirs
        info_.recv_bytes = mtot * 8 * fwdRows + sentCells * (8 * (nSp + 3) + 4);
        info_.n_send_partners = nS; info_.n_recv_partners = (int32_t)recvFrom_.size();
        return LBCHEM_OK;
    }

    std::unique_ptr<Backend> be_;
    std::unique_ptr<Transport> tp_;
    lbchem_lb_options opt_;
    std::string err_;
    lbchem_lb_info info_{};
    bool firstTime_ = true;                 // reference .H:104
    int iter_;
    double totalCost_ = 0.0;                // totalCpuTime_
    long long lastN_ = -1;
    void* lastCost_ = nullptr; long long lastCap_ = 0;      // cellDataList_[i].cpuTime() of the last step
    std::vector<int32_t> sendTo_, recvFrom_; std::vector<double> sendAmt_, gathered_;   // sendAndReceiveData_
    std::vector<long long> bounds_, recvSizes_;
    std::vector<lbchem_p2p_op> ops_;
    void *rY_ = nullptr, *rRR_ = nullptr, *rT_ = nullptr, *rp_ = nullptr, *rrho_ = nullptr, *rdT_ = nullptr, *rdtc_ = nullptr,
         *rTout_ = nullptr, *rcost_ = nullptr, *rstatus_ = nullptr;
    long long rcap_ = 0;
    void *oTout_ = nullptr, *ocost_ = nullptr, *ostatus_ = nullptr, *tdtc_ = nullptr;
    long long oCap_[3] = {0, 0, 0}, tdCap_ = 0;
};

}  // namespace lbcore

// the opaque C handle wraps a Driver
struct lbchem_lb_t {
    std::unique_ptr<lbcore::Driver> drv;
    lbchem_handle chem = nullptr;            // device build: the chemistry handle (not owned)
    void* extra = nullptr;                   // backend-specific staging (host-pointer entry point)
    std::string err;
};
