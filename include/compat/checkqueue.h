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
// reduction as :77, with no exchange between devices.  Exact for OP_CHECKSIG under SCRIPT_VERIFY_NULLFAIL (the
// post-fork flag set); see INTEGRATION.md §2 for the callers that must stay synchronous.
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
    std::vector<std::unique_ptr<bcp::Recorder>> recorders; // one per thread that evaluated checks for this block
    unsigned long nEpoch = 0;       // bumped by Wait(): a thread's recorder is only valid within one block

    bcp::Recorder *NewRecorder() { // called with mtx held
        recorders.emplace_back(new bcp::Recorder());
        return recorders.back().get();
    }

    // Shared by workers and the master.  A worker sleeps when the queue is empty; the master returns once
    // nothing is outstanding.
    bool Run(bool fMaster) {
        std::vector<T> mine;
        mine.reserve(nBatchSize);
        bcp::Recorder *rec = nullptr;
        unsigned long recEpoch = 0;
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
                if (rec == nullptr || recEpoch != nEpoch) {
                    rec = NewRecorder();
                    recEpoch = nEpoch;
                }
            }
            bcp::Recorder::Scope scope(rec);
            for (T &check : mine)
                if (fOk) fOk = check();
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
        Run(true);
        std::vector<std::unique_ptr<bcp::Recorder>> recs;
        bool hostOk;
        {
            std::lock_guard<std::mutex> lock(mtx);
            hostOk = fResult;
            fResult = true;
            recs.swap(recorders); // nothing is outstanding: no thread is inside a check
            nEpoch++;
        }
        if (!hostOk) return false; // the reference would have stopped evaluating too: nothing to launch
        return bcp::VerifyRecorded(recs);
    }

    //! Add a batch of checks; the elements are moved out of vChecks (:147-149).
    void Add(std::vector<T> &vChecks) {
        std::lock_guard<std::mutex> lock(mtx);
        for (T &check : vChecks) {
            pending.emplace_back();
            check.swap(pending.back());
        }
        nOutstanding += (unsigned int)vChecks.size();
        if (vChecks.size() == 1) cvWorker.notify_one();
        else if (vChecks.size() > 1) cvWorker.notify_all();
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
