// include/compat/checkqueue.h — level B1 of the boundary: a drop-in for the node's
// src/checkqueue.h with the same template surface, so that src/validation.cpp compiles unchanged:
//
//   CCheckQueue<T>(unsigned nBatchSize)          src/checkqueue.h:131-133
//   void Thread()                                 :138   worker loop (returns only after Quit())
//   bool Wait()                                   :142   master joins the work, returns the AND, resets the queue
//   void Add(std::vector<T>& vChecks)             :145   moves the checks out of the caller's vector
//   bool IsIdle()                                 :160
//   CCheckQueueControl<T>(CCheckQueue<T>*)        :170-199  RAII: asserts idle, Add, Wait, waits in the destructor;
//                                                          a nullptr queue means "skip the checks" and yields true
//   T needs: default constructor, bool operator()(), void swap(T&)            (:19-20)
//
// What is different inside: host threads still run every T::operator()() — script interpretation, sighash and
// encoding checks stay on the CPU — but the signature checker they reach is in RECORDING mode
// (bcp::Recorder, bcp_recorder.h): it lax-parses the DER signature, appends (sighash, pubkey, r||s) to the
// calling thread's batch and answers true.  Wait() then hands the concatenated batch to the GPUs
// (bcp_gpu_ecdsa_verify_batch with BCP_ECDSA_FLAG_NORMALIZE_S == CPubKey::Verify semantics), sharded by index
// range over every initialised device, and returns host_result AND all(verdicts) — the same `fAllOk &= fOk`
// reduction as :77, with no exchange between devices.
//
// Replay.  "true" is only an assumption, and a script may legitimately see a failing
// comparison: OP_CHECKMULTISIG walks the keys until one matches (src/script/interpreter.cpp:1054-1187), and
// without SCRIPT_VERIFY_NULLFAIL a script may branch on a failed OP_CHECKSIG.  So every check that reached a
// signature comparison is kept until Wait().  A check whose assumptions the devices all confirmed is final with
// the result it returned.  A check with at least one refuted assumption is run again with the verdicts known so
// far answered exactly and only new comparisons assumed; those go to the devices in the next round.  A run is
// final as soon as everything it assumed has been confirmed, and each round settles at least one new comparison
// per open check, so this ends after at most (number of distinct comparisons the script can make) rounds — two
// for an m-of-n multisig signed by keys in order.  No signature is ever verified on the CPU.
// A device error makes Wait() return false (never a silent accept).
//
// Threading primitives are std:: (the reference uses boost::mutex / condition_variable; Boost is not a
// dependency of this project).
#ifndef BCP_COMPAT_CHECKQUEUE_H
#define BCP_COMPAT_CHECKQUEUE_H

#include <algorithm>
#include <cassert>
#include <condition_variable>
#include <memory>
#include <mutex>
#include <thread>
#include <vector>

#include "bcp_recorder.h"

template <typename T> class CCheckQueue {
private:
    std::mutex mtx;
    std::condition_variable cvWorker, cvMaster;
    std::vector<T> pending;         // LIFO: order of evaluation does not matter, only the AND (:38-40)
    unsigned int nOutstanding = 0;  // queued + currently being evaluated
    int nWorkers = 0, nIdleWorkers = 0;
    bool fResult = true;
    bool fQuit = false;
    const unsigned int nBatchSize;
    // One per thread that evaluated checks for this block: its batch, and the checks that reached at least one
    // signature comparison (moved in), each with the end of its range in the batch and the result it returned.
    struct Lane {
        bcp::Recorder rec;
        std::vector<T> kept;
        std::vector<size_t> end;
        std::vector<uint8_t> result;
    };
    std::vector<std::unique_ptr<Lane>> lanes;
    unsigned long nEpoch = 0;       // bumped by Wait(): a thread's lane is only valid within one block

    Lane *NewLane() { // called with mtx held
        lanes.emplace_back(new Lane());
        return lanes.back().get();
    }

    // Settles the assumed verdicts (see "Replay" above).  Runs on the thread that called Wait().
    static bool Resolve(std::vector<std::unique_ptr<Lane>> &ls) {
        std::vector<bcp::Recorder *> recs;
        for (auto &l : ls) recs.push_back(&l->rec);
        std::vector<uint8_t> ok;
        if (!bcp::VerifyRecorders(recs, ok)) return false;
        bcp::KnownVerdicts known;
        std::vector<T *> open;
        size_t base = 0;
        for (auto &l : ls) {
            size_t b = 0;
            for (size_t i = 0; i < l->kept.size(); i++) {
                const size_t e = l->end[i];
                bool confirmed = true;
                for (size_t j = b; j < e; j++) confirmed = confirmed && ok[base + j];
                if (confirmed) {
                    if (!l->result[i]) return false;
                } else {
                    for (size_t j = b; j < e; j++) known[l->rec.keyAt(j)] = ok[base + j] != 0;
                    open.push_back(&l->kept[i]);
                }
                b = e;
            }
            base += l->rec.size();
        }
        while (!open.empty()) {
            bcp::Recorder rr;
            rr.known = &known;
            std::vector<size_t> end;
            std::vector<uint8_t> result;
            {
                bcp::Recorder::Scope scope(&rr);
                for (T *c : open) {
                    const size_t before = rr.size();
                    const bool r = (*c)();
                    if (rr.size() == before && !r) return false; // nothing assumed: final
                    end.push_back(rr.size());
                    result.push_back(r);
                }
            }
            if (rr.size() == 0) return true;
            std::vector<bcp::Recorder *> one{&rr};
            if (!bcp::VerifyRecorders(one, ok)) return false;
            std::vector<T *> next;
            size_t b = 0;
            for (size_t i = 0; i < open.size(); i++) {
                const size_t e = end[i];
                bool confirmed = true;
                for (size_t j = b; j < e; j++) confirmed = confirmed && ok[j];
                if (confirmed) {
                    if (!result[i]) return false;
                } else {
                    for (size_t j = b; j < e; j++) known[rr.keyAt(j)] = ok[j] != 0;
                    next.push_back(open[i]);
                }
                b = e;
            }
            open.swap(next);
        }
        return true;
    }

    // Shared by workers and the master.  A worker sleeps when the queue is empty; the master returns once
    // nothing is outstanding.
    bool Run(bool fMaster) {
        std::vector<T> mine;
        mine.reserve(nBatchSize);
        Lane *lane = nullptr;
        unsigned long laneEpoch = 0;
        bool fRegistered = false;
        unsigned int nDone = 0;
        bool fOk = true;
        for (;;) {
            {
                std::unique_lock<std::mutex> lock(mtx);
                if (nDone) {
                    fResult = fResult && fOk;
                    nOutstanding -= nDone;
                    nDone = 0;
                    if (nOutstanding == 0 && !fMaster) cvMaster.notify_one();
                } else if (!fMaster && !fRegistered) {
                    nWorkers++; // registered in the same critical section in which the worker first goes idle
                    fRegistered = true;
                }
                while (pending.empty()) {
                    if (fMaster && nOutstanding == 0) return true; // caller (Wait) finishes under its own lock
                    if (!fMaster && fQuit) {
                        nWorkers--;
                        return false;
                    }
                    if (fMaster) {
                        cvMaster.wait(lock);
                    } else {
                        nIdleWorkers++;
                        cvWorker.wait(lock);
                        nIdleWorkers--;
                    }
                }
                // share what is queued between the threads that could take it (same rule as :109-111)
                unsigned int share = (unsigned int)pending.size() / (unsigned int)(nWorkers + 1);
                unsigned int take = std::max(1u, std::min(nBatchSize, share));
                mine.resize(take);
                for (unsigned int i = 0; i < take; i++) {
                    mine[i].swap(pending.back());
                    pending.pop_back();
                }
                fOk = fResult; // once a check has failed the rest is only drained, not evaluated (:121-126)
                if (lane == nullptr || laneEpoch != nEpoch) {
                    lane = NewLane();
                    laneEpoch = nEpoch;
                }
            }
            bcp::Recorder::Scope scope(&lane->rec);
            for (T &check : mine) {
                if (!fOk) continue;
                const size_t before = lane->rec.size();
                const bool r = check();
                if (lane->rec.size() == before) {
                    fOk = r; // no signature comparison was reached: the result is final
                } else {     // the result rests on assumed verdicts: settled in Wait()
                    lane->kept.emplace_back();
                    check.swap(lane->kept.back());
                    lane->end.push_back(lane->rec.size());
                    lane->result.push_back(r);
                }
            }
            nDone = (unsigned int)mine.size();
            mine.clear();
        }
    }

public:
    explicit CCheckQueue(unsigned int nBatchSizeIn) : nBatchSize(nBatchSizeIn ? nBatchSizeIn : 1) {}
    ~CCheckQueue() {}

    //! Worker thread body.
    void Thread() { Run(false); }

    //! Wait until all queued checks are evaluated, run the deferred signature batch, return the AND, reset.
    bool Wait() {
        {
            std::lock_guard<std::mutex> lock(mtx);
            if (!pending.empty()) cvWorker.notify_all();
        }
        Run(true);
        std::vector<std::unique_ptr<Lane>> ls;
        bool hostOk;
        {
            std::lock_guard<std::mutex> lock(mtx);
            hostOk = fResult;
            fResult = true;
            ls.swap(lanes); // nothing is outstanding: no thread is inside a check
            nEpoch++;
        }
        if (!hostOk) return false; // the reference would have stopped evaluating too: nothing to launch
        return Resolve(ls);
    }

    //! Add a batch of checks; the elements are moved out of vChecks (:147-149).
    void Add(std::vector<T> &vChecks) {
        std::lock_guard<std::mutex> lock(mtx);
        for (T &check : vChecks) {
            pending.emplace_back();
            check.swap(pending.back());
        }
        nOutstanding += (unsigned int)vChecks.size();
        // Workers are woken only when a batch worth of work has piled up (the reference wakes one per Add, :152-155).
        // With the signature arithmetic deferred a check takes a few microseconds — less than waking a thread — and
        // ConnectBlock adds one transaction at a time; waking per Add made sixteen threads fight over the mutex and
        // slowed the producer 3x.  Nothing is lost by waiting: under a backlog every further Add wakes one more worker,
        // and Wait() wakes everybody and joins the work itself.
        if (pending.size() >= nBatchSize) {
            if (vChecks.size() > nBatchSize) cvWorker.notify_all();
            else cvWorker.notify_one();
        }
    }

    bool IsIdle() {
        std::lock_guard<std::mutex> lock(mtx);
        return nWorkers == nIdleWorkers && nOutstanding == 0 && fResult;
    }

    //! Extension: ask the worker threads to return from Thread() (the reference relies on boost thread
    //! interruption for shutdown, src/init.cpp).
    void Quit() {
        std::lock_guard<std::mutex> lock(mtx);
        fQuit = true;
        cvWorker.notify_all();
    }
};

template <typename T> class CCheckQueueControl {
private:
    CCheckQueue<T> *pqueue;
    bool fDone;

public:
    CCheckQueueControl() = delete;
    CCheckQueueControl(const CCheckQueueControl &) = delete;
    CCheckQueueControl &operator=(const CCheckQueueControl &) = delete;
    explicit CCheckQueueControl(CCheckQueue<T> *pqueueIn) : pqueue(pqueueIn), fDone(false) {
        if (pqueue != nullptr) {
            bool isIdle = pqueue->IsIdle(); // only one block in flight at a time (:176-183)
            assert(isIdle);
            (void)isIdle;
        }
    }
    bool Wait() {
        if (pqueue == nullptr) return true;
        bool fRet = pqueue->Wait();
        fDone = true;
        return fRet;
    }
    void Add(std::vector<T> &vChecks) {
        if (pqueue != nullptr) pqueue->Add(vChecks);
    }
    ~CCheckQueueControl() {
        if (!fDone) Wait();
    }
};

#endif // BCP_COMPAT_CHECKQUEUE_H
