// include/dindel_b200_driver.hpp — the batched window loop for the reference's driver.
//
// DetInDel::detectIndels (DInDel.cpp:1265-1422) handles one window at a time: getReads -> getHaplotypes ->
// computeLikelihoods -> filterHaplotypes -> diploidGLF / Bayes EM -> glf lines.  With the likelihood engine on a
// GPU that loop leaves the device idle (one 0.5 ms call per 400-pair window) and one core busy.  This header splits it:
//
//   prepare   (host, N worker threads)   getReads (:885) + getHaplotypes (:1526) per window, exactly as the reference
//                                        runs them, each worker on its own DetInDel (own BAM / fasta handles, own read
//                                        buffer) over a contiguous run of windows; the window's haplotypes, reads and
//                                        variant entries are copied into ONE dindel_b200::WindowBatch
//   flush     (GPU, one call)            dindel_b200_process: K1 + K4 + K2 for every window of the batch
//   finish    (host, N worker threads)   the reference's own empiricalDistributionMethod (:380) per window with its
//                                        getHaplotypes / computeLikelihoods / filterHaplotypes calls answered from the
//                                        batch (thread-local replay slot), glf lines buffered per window
//   write     (host, in window order)    the buffered lines, byte-identical to the window-at-a-time loop
//
// It is written against the reference's own types and is compiled INTO the reference's DInDel.cpp (INTEGRATION.md
// section 2: one `friend struct dindel_b200_driver;` in class DetInDel, this include, and a three-line forward at the
// top of detectIndels).  Nothing here computes a likelihood: no CPU fallback exists.
//
// Behaviour kept from detectIndels: window index numbering, the sorted-windows check, per-window try/catch ->
// "error_<msg>" glf line (:1369-1408), params.tid per window.  The read buffer (getReads' incremental BAM fetch,
// :928-997) is restarted at the first window of every worker's run and after every skipped window; for one BAM file
// the reads it returns do not depend on where it was restarted (same set, BAM order), which is what makes the runs
// independent.  With several BAM files (--bamFiles) the order of reads inside the buffer depends on the restart
// points, so the driver then prepares with ONE worker and, when a window fails late (an exception raised after the
// likelihoods, e.g. "Could not find indel allele"), re-prepares the windows after it exactly as the sequential loop
// would have (reset buffer).
#ifndef DINDEL_B200_DRIVER_HPP
#define DINDEL_B200_DRIVER_HPP
#include <pthread.h>
#include <time.h>
#include <deque>
#include "dindel_b200.hpp"

namespace dindel_b200 {
// what the patched member functions find in the thread-local slot while a window is being finished
template <class HapT, class ReadT>
struct Replay {
    WindowBatch<HapT, ReadT> *batch;
    int w;                              // window index inside the batch, -1: not on the GPU (skip / too large)
    const std::vector<HapT> *haps;
    uint32_t leftPos, rightPos;
    bool skip;
    Inference inf;
    pthread_mutex_t *gpuLock;           // serialises per-window GPU calls of the finish threads (K3)
    void windowEM(const std::vector<uint8_t> &compatible, size_t nSets, double a0, double tol, std::vector<double> &loglik,
                  std::vector<int32_t> &iters, std::vector<double> &pi, std::vector<double> &margin)
    {
        pthread_mutex_lock(gpuLock);
        try { batch->windowEM(w, compatible, nSets, a0, tol, loglik, iters, pi, margin); }
        catch (...) { pthread_mutex_unlock(gpuLock); throw; }
        pthread_mutex_unlock(gpuLock);
    }
    // K6 for this window (dindel_b200::deviceStats' provider)
    bool statsResident() const { return w >= 0 && batch && batch->statsResident(); }
    void windowStats(const std::vector<StatsJob> &jobs, std::vector<dindel_read_stats> &out)
    {
        pthread_mutex_lock(gpuLock);
        try { batch->windowStats(w, jobs, out); }
        catch (...) { pthread_mutex_unlock(gpuLock); throw; }
        pthread_mutex_unlock(gpuLock);
    }
};
} // namespace dindel_b200

typedef dindel_b200::Replay<Haplotype, Read> dindel_b200_replay_t;
static __thread dindel_b200_replay_t *dindel_b200_replay = 0;

struct dindel_b200_driver {
    typedef dindel_b200::WindowBatch<Haplotype, Read> Batch;
    struct Win {
        AlignedCandidates cand;
        int index;
        uint32_t pos, leftPos, rightPos;   // as read from the window file (DInDel.cpp:1322-1326)
        uint32_t hLeft, hRight;            // after getHaplotypes moved them (:393)
        std::vector<Read> reads;
        std::vector<Haplotype> haps;
        bool skipHaps, failed, prepared, stringThrown;
        std::string message;               // reason of a skipped window
        int bw;                            // index inside the batch (-1: not submitted)
        std::string glf;                   // the window's glf lines
        Win() : index(0), pos(0), leftPos(0), rightPos(0), hLeft(0), hRight(0), skipHaps(false), failed(false), prepared(false), stringThrown(true), bw(-1) {}
    };
    struct Worker {
        dindel_b200_driver *drv;
        DetInDel *det;
        int id;
        size_t first, last;                // its run of windows inside the current batch
        int phase;
        double tPrepare, tFinish, tReads, tHaps, tAdd;   // thread-seconds: phases; getReads / getHaplotypes / batch add
        dindel_b200_replay_t replay;
        Worker() : drv(0), det(0), id(0), first(0), last(0), phase(0), tPrepare(0.0), tFinish(0.0), tReads(0.0), tHaps(0.0), tAdd(0.0) {}
    };

    DetInDel &master;
    // The engine (CUDA context: ~1.4 s) is created on a helper thread while the workers already fetch reads and build
    // haplotypes; a worker parks its prepared windows until the engine is there and adds them to the batch then.
    typedef dindel_b200::Engine *(*EngineFactory)(void *);
    EngineFactory makeEngine;
    void *makeEngineArg;
    dindel_b200::Engine *gpu;
    volatile int engineState;              // 0 = being created, 1 = ready, -1 = failed (engineError)
    std::string engineError;
    pthread_t engineThread;
    bool engineThreadStarted, engineJoined;
    pthread_mutex_t engineLock;
    double tEngine;
    std::string bamArg;
    int multipleFiles;
    std::vector<Worker> workers;
    std::vector<Win> wins;
    Batch *batch;
    pthread_mutex_t addLock, gpuLock;
    bool inferOnGpu;
    double tFlush, tRead, tWrite;
    long nWindows, nBatches, nLateFail;
    dindel_b200::Inference totals;

    dindel_b200_driver(DetInDel &m, EngineFactory f, void *farg, const std::string &bam, int multi, bool infer)
        : master(m), makeEngine(f), makeEngineArg(farg), gpu(0), engineState(0), engineThreadStarted(false), engineJoined(false),
          tEngine(0.0), bamArg(bam), multipleFiles(multi), batch(0), inferOnGpu(infer), tFlush(0.0), tRead(0.0), tWrite(0.0),
          nWindows(0), nBatches(0), nLateFail(0)
    {
        pthread_mutex_init(&addLock, 0);
        pthread_mutex_init(&gpuLock, 0);
        pthread_mutex_init(&engineLock, 0);
    }
    static void *engineMain(void *arg)
    {
        dindel_b200_driver &d = *(dindel_b200_driver *)arg;
        const double t0 = now();
        try {
            d.gpu = d.makeEngine(d.makeEngineArg);
            d.batch = new Batch(*d.gpu, d.master.params.obsParams.maxMismatch, d.master.params.filterHaplotypes, true);
            d.tEngine = now() - t0;
            __sync_synchronize();
            d.engineState = 1;
        } catch (std::string s) {
            d.engineError = s;
            __sync_synchronize();
            d.engineState = -1;
        }
        return 0;
    }
    void startEngine()
    {
        if (pthread_create(&engineThread, 0, engineMain, this) == 0) engineThreadStarted = true;
        else engineMain(this);
    }
    // blocks until the engine exists (or failed); true when it is usable
    bool waitEngine()
    {
        pthread_mutex_lock(&engineLock);
        if (engineThreadStarted && !engineJoined) { pthread_join(engineThread, 0); engineJoined = true; }
        pthread_mutex_unlock(&engineLock);
        __sync_synchronize();
        return engineState == 1;
    }
    ~dindel_b200_driver()
    {
        waitEngine();
        for (size_t k = 1; k < workers.size(); k++) delete workers[k].det;
        delete batch;
    }
    static double now() { timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec; }

    // one DetInDel per worker: worker 0 is the caller's own object
    void makeWorkers(int n)
    {
        workers.resize((size_t)n);
        for (int k = 0; k < n; k++) {
            Worker &wk = workers[k];
            wk.drv = this; wk.id = k;
            if (k == 0) wk.det = &master;
            else {
                wk.det = new DetInDel(bamArg, master.params, multipleFiles);
                wk.det->params = master.params;       // --libFile switches set after construction (DInDel.cpp:4268-4272)
                wk.det->libraries = master.libraries;
            }
            wk.replay.gpuLock = &gpuLock;
            wk.replay.inf.enable(inferOnGpu);
        }
    }

    // ---- prepare: DInDel.cpp:1361-1362 + :393-399 for the worker's run of windows ----
    static void prepareRun(Worker &wk)
    {
        dindel_b200_driver &d = *wk.drv;
        DetInDel &det = *wk.det;
        std::vector<Read *> readBuffer;
        uint32_t oldLeftPos = 0, oldRightFetchReadPos = 0;
        bool reset = true;
        std::string oldTid("-1");
        std::vector<size_t> parked;
        for (size_t i = wk.first; i < wk.last; i++) {
            Win &x = d.wins[i];
            if (x.prepared) continue;
            det.params.tid = x.cand.tid;
            if (det.params.tid != oldTid) { reset = true; oldTid = det.params.tid; oldLeftPos = 0; }
            x.failed = false; x.bw = -1; x.message = "ok";
            try {
                const double tr0 = now();
                try { det.getReads(x.leftPos, x.rightPos, x.reads, oldLeftPos, oldRightFetchReadPos, readBuffer, reset); }
                catch (...) { wk.tReads += now() - tr0; throw; }
                const double tr1 = now();
                wk.tReads += tr1 - tr0;
                if (det.params.inferenceMethod != "empirical") throw std::string("Unknown inference method");
                x.hLeft = x.leftPos; x.hRight = x.rightPos;
                x.haps.clear();
                try { x.skipHaps = det.getHaplotypes(x.haps, x.reads, x.pos, x.hLeft, x.hRight, x.cand); }
                catch (...) { wk.tHaps += now() - tr1; throw; }
                wk.tHaps += now() - tr1;
                // windows the reference refuses after getHaplotypes (:395-399) or skips (:403) never reach the GPU
                const bool tooLarge = int(x.reads.size() * x.haps.size()) > det.params.maxHapReadProd; // thrown at :395-399
                if (!x.skipHaps && !tooLarge && !x.haps.empty()) parked.push_back(i);
                if (d.engineState == 1) { addParked(d, parked); }
                reset = tooLarge; // a skipped window restarts the read buffer (:1405)
            } catch (std::string s) { x.failed = true; x.message = s; x.stringThrown = true; reset = true; }
            catch (std::bad_alloc) { x.failed = true; x.message = "bad_alloc"; x.stringThrown = false; reset = true; }
            catch (std::exception &e) { x.failed = true; x.message = std::string("exception_").append(e.what()); x.stringThrown = false; reset = true; }
            x.prepared = true;
            oldLeftPos = x.leftPos;
        }
        for (size_t r = 0; r < readBuffer.size(); r++) if (readBuffer[r] != NULL) delete readBuffer[r];
        if (d.waitEngine()) addParked(d, parked);
        else for (size_t k = 0; k < parked.size(); k++) { // no device: what the engine's constructor threw, window by window
            Win &x = d.wins[parked[k]];
            x.failed = true; x.stringThrown = true; x.message = d.engineError;
        }
    }
    // the prepared windows a worker parked while the engine was being created -> the batch
    static void addParked(dindel_b200_driver &d, std::vector<size_t> &parked)
    {
        for (size_t k = 0; k < parked.size(); k++) {
            Win &x = d.wins[parked[k]];
            pthread_mutex_lock(&d.addLock);
            try { x.bw = d.batch->add(x.haps, x.reads, x.hLeft); }
            catch (std::string s) { x.failed = true; x.stringThrown = true; x.message = s; }
            catch (...) { pthread_mutex_unlock(&d.addLock); throw; }
            pthread_mutex_unlock(&d.addLock);
        }
        parked.clear();
    }

    // the "error_<msg>" line of a skipped window (DInDel.cpp:1389-1402)
    static void errorLine(DetInDel &det, OutputData &glfData, const Win &x, std::string message, bool thrownString)
    {
        if (thrownString) for (size_t k = 0; k < message.size(); k++) if (message[k] == ' ') message[k] = '_'; // :1370
        message = std::string("error_").append(message);
        std::cerr << "skipped " << det.params.tid << " " << x.pos << " reason: " << message << std::endl;
        OutputData::Line gline(glfData);
        gline.set("msg", message);
        gline.set("index", x.index);
        gline.set("tid", det.params.tid);
        gline.set("lpos", x.leftPos);
        gline.set("rpos", x.rightPos);
        glfData.output(gline);
    }

    // ---- finish: DInDel.cpp:1365-1408 for the worker's run, the GPU's results replayed into the reference's calls ----
    static void finishRun(Worker &wk)
    {
        dindel_b200_driver &d = *wk.drv;
        DetInDel &det = *wk.det;
        std::ofstream nullOut;
        OutputData oData = det.params.makeOutputData(nullOut);
        std::ostringstream os;
        OutputData glfData = det.params.makeGLFOutputData(os);
        for (size_t i = wk.first; i < wk.last; i++) {
            Win &x = d.wins[i];
            os.str(std::string()); os.clear();
            det.params.tid = x.cand.tid;
            if (x.failed) {
                errorLine(det, glfData, x, x.message, x.stringThrown);
                x.glf = os.str();
                continue;
            }
            wk.replay.batch = d.batch; wk.replay.w = x.bw; wk.replay.haps = &x.haps;
            wk.replay.leftPos = x.hLeft; wk.replay.rightPos = x.hRight; wk.replay.skip = x.skipHaps;
            dindel_b200_replay = &wk.replay;
            bool skipped = false, str = true;
            std::string message;
            try {
                det.empiricalDistributionMethod(x.index, x.reads, x.pos, x.leftPos, x.rightPos, x.cand, oData, glfData);
            } catch (std::string s) { skipped = true; message = s; }
            catch (std::bad_alloc) { skipped = true; str = false; message = "bad_alloc"; }
            catch (std::exception &e) { skipped = true; str = false; message = std::string("exception_").append(e.what()); }
            dindel_b200_replay = 0;
            if (skipped) {
                x.failed = true; x.message = message;
                errorLine(det, glfData, x, message, str);
            }
            x.glf = os.str();
        }
    }

    static void *threadMain(void *arg)
    {
        Worker &wk = *(Worker *)arg;
        const double t0 = now();
        if (wk.phase == 0) { prepareRun(wk); wk.tPrepare += now() - t0; }
        else { finishRun(wk); wk.tFinish += now() - t0; }
        return 0;
    }

    void runPhase(int phase, size_t first, size_t last)
    {
        const size_t n = last - first, nt = workers.size();
        std::vector<pthread_t> th(nt);
        for (size_t k = 0; k < nt; k++) {
            Worker &wk = workers[k];
            wk.phase = phase;
            wk.first = first + n * k / nt; wk.last = first + n * (k + 1) / nt;
        }
        // the last non-empty run on the calling thread, the others on their own threads
        std::vector<int> started(nt, 0);
        int mine = -1;
        for (size_t k = 0; k < nt; k++) if (workers[k].first != workers[k].last) mine = (int)k;
        for (size_t k = 0; k < nt; k++) {
            if ((int)k == mine || workers[k].first == workers[k].last) continue;
            if (pthread_create(&th[k], 0, threadMain, &workers[k]) != 0) threadMain(&workers[k]);
            else started[k] = 1;
        }
        if (mine >= 0) threadMain(&workers[mine]);
        for (size_t k = 0; k < nt; k++) if (started[k]) pthread_join(th[k], 0);
    }

    // DetInDel::detectIndels (DInDel.cpp:1265-1422), batched
    void run(const std::string &variantsFileName, int nThreads, size_t batchWindows)
    {
        std::ofstream glfOutput;
        std::string glfFile = master.params.fileName; glfFile.append(".glf.txt");
        glfOutput.open(glfFile.c_str());
        if (!glfOutput.is_open()) throw(std::string("Cannot open file ").append(glfFile).append(" for writing."));
        OutputData glfData = master.params.makeGLFOutputData(glfOutput);
        glfData.outputLine(glfData.headerString());
        VariantFile vf(variantsFileName);
        const bool oneBam = master.myBams.size() == 1;
        startEngine();
        makeWorkers(oneBam ? nThreads : 1);
        int index = 0;
        uint32_t oldLeftPos = 0;
        std::string oldTid("-1");
        std::deque<Win> carry;              // windows to prepare again after a late failure (several BAM files only)
        bool eof = false;
        while (!eof || !carry.empty()) {
            // ---- read the next batch of window lines (DInDel.cpp:1310-1345) ----
            double t0 = now();
            wins.clear();
            while (!carry.empty() && wins.size() < batchWindows) { wins.push_back(carry.front()); carry.pop_front(); }
            while (wins.size() < batchWindows && !(eof = vf.eof())) {
                AlignedCandidates cand = vf.getLineVector(master.params.varFileIsOneBased);
                if (cand.variants.size() == 0) continue;
                Win x;
                x.cand = cand;
                x.leftPos = cand.leftPos; x.rightPos = cand.rightPos; x.pos = cand.centerPos;
                if (cand.tid != oldTid) { oldTid = cand.tid; oldLeftPos = 0; }
                if (x.leftPos < oldLeftPos) {
                    std::cerr << "leftPos: " << x.leftPos << " oldLeftPos: " << oldLeftPos << std::endl;
                    std::cerr << "Candidate variant files must be sorted on left position of window!" << std::endl;
                    exit(1);
                }
                x.index = ++index;
                oldLeftPos = x.leftPos;
                wins.push_back(x);
            }
            tRead += now() - t0;
            if (wins.empty()) break;
            nBatches++;
            if (nBatches > 1 && waitEngine()) batch->clear();
            runPhase(0, 0, wins.size());
            t0 = now();
            if (waitEngine()) batch->flush();
            tFlush += now() - t0;
            runPhase(1, 0, wins.size());
            // ---- write in window order; several BAM files: a late failure restarts the read buffer for what follows ----
            t0 = now();
            size_t upto = wins.size();
            if (!oneBam)
                for (size_t i = 0; i + 1 < wins.size(); i++)
                    if (wins[i].failed && wins[i].prepared && wins[i].bw >= 0) { upto = i + 1; nLateFail++; break; }
            for (size_t i = 0; i < upto; i++) glfOutput << wins[i].glf;
            nWindows += (long)upto;
            for (size_t i = upto; i < wins.size(); i++) {
                Win x = wins[i];
                x.prepared = false; x.failed = false; x.bw = -1; x.reads.clear(); x.haps.clear(); x.glf.clear();
                carry.push_back(x);
            }
            tWrite += now() - t0;
        }
        glfOutput.close();
        for (size_t k = 0; k < workers.size(); k++) {
            totals.nWindows += workers[k].replay.inf.nWindows; totals.nRecheck += workers[k].replay.inf.nRecheck;
            totals.nEmSets += workers[k].replay.inf.nEmSets; totals.nEmMismatch += workers[k].replay.inf.nEmMismatch;
            totals.nEmUnsupported += workers[k].replay.inf.nEmUnsupported;
            for (int i = 0; i < 6; i++) totals.nTrip[i] += workers[k].replay.inf.nTrip[i];
        }
        double tp = 0.0, tf = 0.0, trd = 0.0, thp = 0.0;
        for (size_t k = 0; k < workers.size(); k++) { tp += workers[k].tPrepare; tf += workers[k].tFinish; trd += workers[k].tReads; thp += workers[k].tHaps; }
        fprintf(stderr, "dindel_b200 host stages (thread-seconds): getReads %.3f getHaplotypes %.3f (of prepare %.3f, the rest waits for the engine / adds to the batch)\n", trd, thp, tp);
        fprintf(stderr, "dindel_b200 batched: windows %ld batches %ld threads %d late_failures %ld; wall: read %.3f s, flush(GPU) %.3f s, write %.3f s; "
                        "thread-seconds: prepare %.3f finish %.3f; engine created in %.3f s (under the first prepare)\n", nWindows, nBatches, (int)workers.size(), nLateFail, tRead, tFlush, tWrite, tp, tf, tEngine);
        if (totals.nWindows)
            fprintf(stderr, "dindel_b200 inference: windows %ld rechecked_on_host %ld em_sets %ld em_set_mismatch %ld em_unsupported_windows %ld\n",
                    totals.nWindows, totals.nRecheck, totals.nEmSets, totals.nEmMismatch, totals.nEmUnsupported);
        if (totals.nRecheck)
            fprintf(stderr, "dindel_b200 guard band, first guard that tripped per rechecked window: compare %ld compare_lin %ld printed %ld printed_lin %ld phred %ld em_stop_margin %ld\n",
                    totals.nTrip[0], totals.nTrip[1], totals.nTrip[2], totals.nTrip[3], totals.nTrip[4], totals.nTrip[5]);
        if (gpu && gpu->nStatsCalls)
            { fprintf(stderr, "dindel_b200 read stats: device_calls %ld jobs %ld\n", gpu->nStatsCalls, gpu->nStatsJobs); gpu->nStatsCalls = 0; }
    }
};
#endif
