// Host runtime behind bert_encode_batch / bert_forward_batch (reference bert.cpp:1374-1444): persistent tokenizer
// threads -> deterministic chunks of sentences -> token-budget micro-batches -> ONE queue that persistent per-GPU
// worker threads pull from (dynamic longest-first list scheduling, no collective, no barrier between chunks).
// Per GPU: three pinned staging slots and three streams, so that the H2D copy of micro-batch k+1 and the D2H copy of
// k-1 overlap the forward of k.  Nothing here depends on timing: which sentences share a micro-batch is a function
// of the token counts only, so results are bit-identical from run to run and for any number of host threads.
#include <algorithm>
#include <chrono>
#include <new>
#include <string>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <numeric>

#include "engine.h"

namespace b200 {

static double now_us() {
    return std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

void Engine::trace(const char * what, long a, long b) {
    if (!tracing) return;
    char buf[160];
    snprintf(buf, sizeof(buf), "%s %ld %ld", what, a, b);
    std::lock_guard<std::mutex> g(trace_mu);
    trace_ev.emplace_back(now_us(), buf);
}

// ------------------------------------------------------------------------------------ HostPool
HostPool::~HostPool() {
    {
        std::lock_guard<std::mutex> g(m_);
        stop_ = true;
    }
    cv_.notify_all();
    for (auto & t : th_) t.join();
}

void HostPool::body(int idx, uint64_t seen) {
    std::unique_lock<std::mutex> lk(m_);
    for (;;) {
        cv_.wait(lk, [&] { return stop_ || gen_ != seen; });
        if (stop_) return;
        seen = gen_;
        if (idx >= want_) continue;
        std::function<void(int)> f = fn_;
        lk.unlock();
        f(idx);
        lk.lock();
        if (--active_ == 0) done_cv_.notify_all();
    }
}

void HostPool::run(int n, std::function<void(int)> fn) {
    std::unique_lock<std::mutex> lk(m_);
    done_cv_.wait(lk, [&] { return active_ == 0; });
    while ((int) th_.size() < n) {
        const int idx = (int) th_.size();
        th_.emplace_back(&HostPool::body, this, idx, gen_);
    }
    fn_ = std::move(fn);
    want_ = n;
    active_ = n;
    ++gen_;
    lk.unlock();
    cv_.notify_all();
}

void HostPool::wait() {
    std::unique_lock<std::mutex> lk(m_);
    done_cv_.wait(lk, [&] { return active_ == 0; });
}

// ------------------------------------------------------------------------------------ micro-batches
int64_t Engine::chunk_target(int chunk_no, int64_t nd) const {
    return nd * std::min<int64_t>(t_cap, (int64_t) 32768 << std::min(chunk_no, 12));
}

// A chunk ends once it holds `target` tokens -- unless what is left of the job (estimated from the mean sentence length
// seen so far: a function of the token counts only) is less than half a chunk, in which case the tail rides along
// instead of becoming a small micro-batch of its own at the end of the call.
bool Engine::chunk_complete(int64_t chunk_tokens, int64_t target, int64_t seen_tokens, int64_t seen_sent, int64_t total_sent) const {
    if (chunk_tokens < target) return false;
    const int64_t rest = seen_sent > 0 ? seen_tokens * (total_sent - seen_sent) / seen_sent : 0;
    return rest >= target / 2;
}

int Engine::piece_budget(int64_t tokens, int64_t nd) const {
    if (nd < 1) nd = 1;
    const int64_t min_piece = std::min<int64_t>(t_cap, std::max<int64_t>(hp.n_max_tokens, 4096));
    int64_t b = (tokens + nd - 1) / nd;
    if (b < min_piece) b = min_piece;
    if (b > t_cap) b = t_cap;
    return (int) b;
}

void Engine::cut_tasks(const std::vector<int> & order, const int32_t * const * tok, const int32_t * len, float * const * out,
                       int budget, const std::shared_ptr<void> & keep, JobState * job, std::vector<MbTask> & tasks) const {
    MbTask cur;
    auto flush = [&]() {
        if (cur.tok.empty()) return;
        cur.keep = keep; cur.job = job;
        tasks.push_back(std::move(cur));
        cur = MbTask();
    };
    for (int idx : order) {
        const int l = len[idx];
        if (!cur.tok.empty() && (cur.tokens + l > budget || (int) cur.tok.size() + 1 > b_cap)) flush();
        cur.tok.push_back(tok[idx]); cur.len.push_back(l); cur.out.push_back(out[idx]);
        cur.tokens += l; cur.max_len = std::max(cur.max_len, l);
    }
    flush();
}

// ------------------------------------------------------------------------------------ one slot of one GPU
namespace {
struct SlotCb { Engine * e; Replica * r; int s; };
// runs on a CUDA driver thread when the D2H copy of a slot has finished: flag the slot and wake the worker
void CUDART_CB slot_done_cb(void * p) {
    SlotCb * cb = (SlotCb *) p;
    {
        std::lock_guard<std::mutex> g(cb->e->q_mu);
        cb->r->slot_ready[cb->s] = 1;
    }
    cb->e->q_cv.notify_all();
}
}  // namespace

bool Engine::slot_submit(Replica & r, int s, MbTask & t, bool notify) {
    trace("submit_begin dev/slot", r.device, s);
    int32_t * ids = r.h_ids[s];
    int32_t * cu = r.h_cu[s];
    const int n = (int) t.tok.size();
    int off = 0;
    for (int i = 0; i < n; ++i) {
        cu[i] = off;
        memcpy(ids + off, t.tok[i], (size_t) t.len[i] * sizeof(int32_t));
        off += t.len[i];
    }
    cu[n] = off;
    const int n_spans = pack_spans(cu, n, cu + n + 1);       // short sentences share attention work items (kernels.cuh)
    const int H = hp.n_embd;
    bool ok = cuda_ok(cudaMemcpyAsync(r.d_ids[s], ids, (size_t) off * 4, cudaMemcpyHostToDevice, r.copy_in), "H2D ids");
    ok = ok && cuda_ok(cudaMemcpyAsync(r.d_cu[s], cu, (size_t) (n + 1 + n_spans + 1) * 4, cudaMemcpyHostToDevice, r.copy_in), "H2D cu");
    ok = ok && cuda_ok(cudaEventRecord(r.h2d_done[s], r.copy_in), "event record");
    ok = ok && cuda_ok(cudaStreamWaitEvent(r.stream, r.h2d_done[s], 0), "stream wait");
    ok = ok && forward(r, r.d_ids[s], r.d_cu[s], n, off, t.max_len, r.d_out[s], n_spans);
    ok = ok && cuda_ok(cudaEventRecord(r.fwd_done[s], r.stream), "event record");
    ok = ok && cuda_ok(cudaStreamWaitEvent(r.copy_out, r.fwd_done[s], 0), "stream wait");
    ok = ok && cuda_ok(cudaMemcpyAsync(r.h_out[s], r.d_out[s], (size_t) n * H * 4, cudaMemcpyDeviceToHost, r.copy_out), "D2H out");
    ok = ok && cuda_ok(cudaEventRecord(r.slot_done[s], r.copy_out), "event record");
    if (ok && notify) {
        static_assert(sizeof(SlotCb) <= sizeof(r.cb_store[0]), "callback storage");
        SlotCb * cb = new (r.cb_store[s]) SlotCb{this, &r, s};
        ok = cuda_ok(cudaLaunchHostFunc(r.copy_out, slot_done_cb, cb), "host callback");
    }
    trace("submit_end tokens/sentences", t.tokens, (long) t.tok.size());
    return ok;
}

void Engine::slot_retire(Replica & r, int s, MbTask & t, bool ok) {
    trace("retire_begin dev/slot", r.device, s);
    if (ok) ok = cuda_ok(cudaEventSynchronize(r.slot_done[s]), "slot sync");
    if (!ok) {                                                 // leave this micro-batch's outputs untouched; drain the GPU
        cudaStreamSynchronize(r.copy_in); cudaStreamSynchronize(r.stream); cudaStreamSynchronize(r.copy_out);
        cudaGetLastError();
    } else {
        const size_t H = (size_t) hp.n_embd;
        const int n = (int) t.out.size();
        const float * src = r.h_out[s];
        for (int i = 0; i < n;) {                              // rows that are adjacent in the caller's memory go in one copy
            int j = i + 1;
            while (j < n && t.out[j] == t.out[j - 1] + H) ++j;
            memcpy(t.out[i], src + (size_t) i * H, (size_t) (j - i) * H * sizeof(float));
            i = j;
        }
    }
    JobState * job = t.job;
    const int tokens = t.tokens;
    t = MbTask();                                              // drops the token storage
    trace("retire_end tokens", tokens);
    {
        std::lock_guard<std::mutex> g(job->m);
        job->pending -= 1;
        job->retired_tokens += tokens;
        if (!ok) job->ok = false;
    }
    job->cv.notify_all();
}

// Latency path: one micro-batch on the calling thread, everything on the compute stream (no events, no hand-offs).
bool Engine::run_inline(Replica & r, MbTask & t) {
    if (!cuda_ok(cudaSetDevice(r.device), "cudaSetDevice")) return false;
    int32_t * ids = r.h_ids[0];
    int32_t * cu = r.h_cu[0];
    const int n = (int) t.tok.size();
    int off = 0;
    for (int i = 0; i < n; ++i) {
        cu[i] = off;
        memcpy(ids + off, t.tok[i], (size_t) t.len[i] * sizeof(int32_t));
        off += t.len[i];
    }
    cu[n] = off;
    const int n_spans = pack_spans(cu, n, cu + n + 1);
    const size_t H = (size_t) hp.n_embd;
    bool ok = cuda_ok(cudaMemcpyAsync(r.d_ids[0], ids, (size_t) off * 4, cudaMemcpyHostToDevice, r.stream), "H2D ids");
    ok = ok && cuda_ok(cudaMemcpyAsync(r.d_cu[0], cu, (size_t) (n + 1 + n_spans + 1) * 4, cudaMemcpyHostToDevice, r.stream), "H2D cu");
    ok = ok && forward(r, r.d_ids[0], r.d_cu[0], n, off, t.max_len, r.d_out[0], n_spans);
    ok = ok && cuda_ok(cudaMemcpyAsync(r.h_out[0], r.d_out[0], (size_t) n * H * 4, cudaMemcpyDeviceToHost, r.stream), "D2H out");
    ok = cuda_ok(cudaStreamSynchronize(r.stream), "stream sync") && ok;
    if (!ok) { cudaGetLastError(); return false; }
    for (int i = 0; i < n; ++i) memcpy(t.out[i], r.h_out[0] + (size_t) i * H, H * sizeof(float));
    return true;
}

// ------------------------------------------------------------------------------------ per-GPU worker threads
void Engine::start_workers() {
    if (!workers.empty()) return;
    stopping = false;
    busy_tokens.assign(reps.size(), 0);
    for (size_t i = 0; i < reps.size(); ++i) workers.emplace_back(&Engine::gpu_worker, this, (int) i);
}

void Engine::stop_workers() {
    {
        std::lock_guard<std::mutex> g(q_mu);
        stopping = true;
    }
    q_cv.notify_all();
    for (auto & t : workers) t.join();
    workers.clear();
}

void Engine::gpu_worker(int ri) {
    Replica & r = reps[ri];
    cudaSetDevice(r.device);
    MbTask inflight[Replica::kSlots];
    bool okf[Replica::kSlots] = {};
    int head = 0, tail = 0, nbusy = 0;
    // A worker only takes the next micro-batch while no other GPU holds fewer tokens in flight: with a handful of large
    // micro-batches per call (two GPUs, one 512 x 512 batch each) a greedy worker would fill its three slots with what the
    // GPU next to it should have run (measured: one call over 2 GPUs at 0.82 of two single-GPU calls before this rule).
    auto fair = [&]() {
        for (size_t j = 0; j < busy_tokens.size(); ++j) if (busy_tokens[j] < busy_tokens[ri]) return false;
        return true;
    };
    for (;;) {
        MbTask t;
        bool got = false, retire = false;
        {
            std::unique_lock<std::mutex> lk(q_mu);
            // sleep until there is a task a free slot can take, or the oldest slot's results have reached the host
            q_cv.wait(lk, [&] {
                return (!queue.empty() && nbusy < Replica::kSlots && fair()) || (nbusy > 0 && (!okf[tail] || r.slot_ready[tail])) ||
                       (stopping && nbusy == 0);
            });
            const bool tail_ready = nbusy > 0 && (!okf[tail] || r.slot_ready[tail]);
            const bool can_take = !queue.empty() && nbusy < Replica::kSlots && fair();
            if (tail_ready && !(can_take && nbusy < 2)) {
                retire = true;                                 // results first (the caller waits for them), unless the GPU is about to run dry
            } else if (can_take) {
                t = std::move(queue.front());
                queue.pop_front();
                r.slot_ready[head] = 0;
                busy_tokens[ri] += t.tokens;
                got = true;
            } else {
                return;                                        // stopping and idle
            }
        }
        if (retire) {
            const int64_t tokens = inflight[tail].tokens;
            slot_retire(r, tail, inflight[tail], okf[tail]);
            tail = (tail + 1) % Replica::kSlots;
            --nbusy;
            {
                std::lock_guard<std::mutex> g(q_mu);
                busy_tokens[ri] -= tokens;
            }
            q_cv.notify_all();                                 // another GPU may now be the one that holds more
        } else if (got) {
            q_cv.notify_all();
            okf[head] = slot_submit(r, head, t, true);
            inflight[head] = std::move(t);
            head = (head + 1) % Replica::kSlots;
            ++nbusy;
        }
    }
}

void Engine::submit(std::vector<MbTask> & tasks, JobState & job) {
    if (tasks.empty()) return;
    {
        std::lock_guard<std::mutex> g(job.m);
        job.pending += (int) tasks.size();
    }
    {
        std::lock_guard<std::mutex> g(q_mu);
        for (MbTask & t : tasks) queue.push_back(std::move(t));
    }
    tasks.clear();
    q_cv.notify_all();
}

void Engine::wait_job(JobState & job) {
    std::unique_lock<std::mutex> lk(job.m);
    job.cv.wait(lk, [&] { return job.pending == 0; });
}

// ------------------------------------------------------------------------------------ bert_forward_batch
bool Engine::run_tokens(const int32_t * const * tokens, const int32_t * n_tokens, int n, float * const * out) {
    if (n <= 0) return true;
    if (reps.empty()) {
        fprintf(stderr, "libbert(b200): this context has no GPU replica (host-only load); there is no CPU path\n");
        return false;
    }
    int64_t total = 0;
    for (int i = 0; i < n; ++i) {
        if (n_tokens[i] > hp.n_max_tokens) {                  // reference bert.cpp:867-871
            fprintf(stderr, "Too many tokens, maximum is %d\n", hp.n_max_tokens);
            return false;
        }
        if (n_tokens[i] <= 0) {
            fprintf(stderr, "bert_forward_batch: input %d has no tokens\n", i);
            return false;
        }
        total += n_tokens[i];
    }
    std::vector<int> order(n);
    std::iota(order.begin(), order.end(), 0);
    std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return n_tokens[a] > n_tokens[b]; });
    JobState job;
    std::vector<MbTask> tasks;
    cut_tasks(order, tokens, n_tokens, out, piece_budget(total, (int64_t) reps.size()), nullptr, &job, tasks);
    if (tasks.size() == 1 && reps.size() == 1) return run_inline(reps[0], tasks[0]);
    submit(tasks, job);                                        // longest sentences first = most expensive first
    wait_job(job);
    return job.ok;
}

// The same chunking and cutting as encode_texts below, on token counts alone (tests / bertx_plan).
int Engine::plan_encode(const int32_t * n_tokens, int n, int nd, int32_t * mb_of_sentence, std::vector<int> & mb_tokens) const {
    mb_tokens.clear();
    std::vector<int> order;
    std::vector<const int32_t *> tok;
    std::vector<float *> out;
    int pos = 0, chunk_no = 0;
    int64_t seen_tokens = 0;
    const int first_kept = n <= kTokBlock ? n : 0;             // latency path: one group, run_tokens' budget
    while (pos < n) {
        const int64_t target = first_kept ? INT64_MAX : chunk_target(chunk_no, nd);
        int64_t chunk_tokens = 0;
        const int first = pos;
        while (pos < n) {
            const int hi = std::min(pos + kTokBlock, n);
            for (; pos < hi; ++pos) { chunk_tokens += n_tokens[pos]; seen_tokens += n_tokens[pos]; }
            if (chunk_complete(chunk_tokens, target, seen_tokens, pos, n)) break;
        }
        const int cnt = pos - first;
        order.resize(cnt);
        std::iota(order.begin(), order.end(), 0);
        std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return n_tokens[first + a] > n_tokens[first + b]; });
        tok.assign(cnt, nullptr); out.assign(cnt, nullptr);
        std::vector<MbTask> tasks;
        // out pointers double as sentence indices here
        for (int i = 0; i < cnt; ++i) out[i] = reinterpret_cast<float *>((uintptr_t) (first + i + 1));
        cut_tasks(order, tok.data(), n_tokens + first, out.data(), piece_budget(chunk_tokens, nd), nullptr, nullptr, tasks);
        for (const MbTask & t : tasks) {
            for (float * o : t.out) mb_of_sentence[(int) ((uintptr_t) o) - 1] = (int32_t) mb_tokens.size();
            mb_tokens.push_back(t.tokens);
        }
        ++chunk_no;
    }
    return (int) mb_tokens.size();
}

// ------------------------------------------------------------------------------------ bert_encode_batch
namespace {
struct TokBlock {
    std::vector<int32_t> ids, lens;
};
struct ChunkStore {                                            // token storage of one chunk; dies with its last micro-batch
    std::vector<TokBlock> blocks;
};
}  // namespace

bool Engine::encode_texts(int n_threads, int n, const char ** texts, float ** out) {
    const int N = hp.n_max_tokens;
    constexpr int kBlock = kTokBlock;
    if (n <= kBlock) {                                         // latency path: tokenise on the calling thread
        std::vector<int32_t> ids((size_t) n * N), lens(n);
        std::vector<const int32_t *> ptrs(n);
        for (int i = 0; i < n; ++i) {
            ptrs[i] = ids.data() + (size_t) i * N;
            lens[i] = tok.tokenize(texts[i], ids.data() + (size_t) i * N, N);
        }
        return run_tokens(ptrs.data(), lens.data(), n, out);
    }
    if (reps.empty()) {
        fprintf(stderr, "libbert(b200): this context has no GPU replica (host-only load); there is no CPU path\n");
        return false;
    }
    if (tracing) { std::lock_guard<std::mutex> g(trace_mu); trace_ev.clear(); }
    trace("encode_begin sentences", n);
    // Tokenizer work items: `sub` sentences each.  Chunks still end on kBlock-sentence boundaries (so the plan stays a
    // function of the token counts alone), but long texts are handed out in smaller pieces: with 16 sentences of 512 tokens
    // per item the first chunk of 64 sentences kept 4 of the 16 threads busy for 1.7 ms before the GPU saw any work.
    int sub = kBlock;
    {
        size_t bytes = 0;
        const int probe = std::min(n, 64);
        for (int i = 0; i < probe; ++i) bytes += strlen(texts[i]);
        const size_t mean = bytes / (size_t) probe;
        if (mean >= 1024) sub = 2; else if (mean >= 256) sub = 4;
    }
    const int per_block = kBlock / sub;                       // work items per chunk-boundary block
    const int nblocks = (n + sub - 1) / sub;
    int nt = n_threads > 0 ? n_threads : 1;
    const int hw = (int) std::thread::hardware_concurrency();
    if (hw > 0 && nt > hw) nt = hw;
    if (nt > nblocks) nt = nblocks;
    const int64_t nd = (int64_t) reps.size();
    const int64_t window = 4 * nd * (int64_t) t_cap;           // tokenised-but-not-finished tokens the workers may run ahead

    JobState job;
    std::vector<TokBlock> blocks(nblocks);
    std::vector<char> ready(nblocks, 0);                       // guarded by job.m
    std::atomic<int> next_block{0};
    int consumer_pos = 0;                                      // guarded by job.m
    int64_t produced = 0;                                      // guarded by job.m
    bool abort = false;                                        // guarded by job.m

    trace("pool_run threads", nt);
    pool.run(nt, [&](int) {
        std::vector<int32_t> scratch((size_t) sub * N);
        int32_t lens[kBlock];
        for (;;) {
            const int b = next_block.fetch_add(1, std::memory_order_relaxed);
            if (b >= nblocks) return;
            {
                std::unique_lock<std::mutex> lk(job.m);
                // the blocks the consumer is about to need always go ahead; later ones respect the window
                job.cv.wait(lk, [&] { return abort || b < consumer_pos + 2 * nt || produced - job.retired_tokens <= window; });
                if (abort) return;
            }
            const int lo = b * sub, hi = std::min(lo + sub, n);
            size_t off = 0;
            bool failed = false;
            try {
                for (int i = lo; i < hi; ++i) {
                    lens[i - lo] = tok.tokenize(texts[i], scratch.data() + off, N);
                    off += (size_t) lens[i - lo];
                }
                blocks[b].ids.assign(scratch.begin(), scratch.begin() + off);
                blocks[b].lens.assign(lens, lens + (hi - lo));
            } catch (...) {
                failed = true;
            }
            {
                std::lock_guard<std::mutex> lk(job.m);
                if (failed) { abort = true; job.ok = false; }
                ready[b] = 1;
                produced += (int64_t) off;
            }
            job.cv.notify_all();
        }
    });
    // whatever happens below (including an exception), the pool threads and the submitted micro-batches reference
    // this frame: stop the former, wait for the latter
    struct Guard {
        Engine & e; JobState & job; bool & abort;
        ~Guard() {
            {
                std::lock_guard<std::mutex> lk(job.m);
                abort = true;
            }
            job.cv.notify_all();
            e.pool.wait();
            e.wait_job(job);
        }
    } guard{*this, job, abort};

    std::vector<MbTask> tasks;
    std::vector<const int32_t *> ptrs;
    std::vector<int32_t> lens;
    std::vector<float *> outs;
    std::vector<int> order;
    int pos = 0, chunk_no = 0;
    int64_t seen_tokens = 0;
    while (pos < nblocks) {
        // deterministic ramp: small first chunks start the GPUs early, then full token-budget chunks
        const int64_t target = chunk_target(chunk_no, nd);
        auto store = std::make_shared<ChunkStore>();
        int64_t chunk_tokens = 0;
        const int first_block = pos;
        while (pos < nblocks) {
            {
                std::unique_lock<std::mutex> lk(job.m);
                job.cv.wait(lk, [&] { return ready[pos] != 0; });
                if (abort) return false;
                consumer_pos = pos + 1;
            }
            job.cv.notify_all();
            chunk_tokens += (int64_t) blocks[pos].ids.size();
            seen_tokens += (int64_t) blocks[pos].ids.size();
            store->blocks.push_back(std::move(blocks[pos]));
            ++pos;
            if (pos % per_block != 0 && pos < nblocks) continue;   // chunks end on kBlock-sentence boundaries only
            if (chunk_complete(chunk_tokens, target, seen_tokens, std::min<int64_t>((int64_t) pos * sub, n), n)) break;
        }
        ptrs.clear(); lens.clear(); outs.clear();
        int sent = first_block * sub;
        for (const TokBlock & tb : store->blocks) {
            size_t off = 0;
            for (int32_t l : tb.lens) {
                ptrs.push_back(tb.ids.data() + off); lens.push_back(l); outs.push_back(out[sent++]);
                off += (size_t) l;
            }
        }
        order.resize(ptrs.size());
        std::iota(order.begin(), order.end(), 0);
        std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return lens[a] > lens[b]; });
        cut_tasks(order, ptrs.data(), lens.data(), outs.data(), piece_budget(chunk_tokens, nd), store, &job, tasks);
        trace("chunk_ready tokens/microbatches", (long) chunk_tokens, (long) tasks.size());
        submit(tasks, job);
        ++chunk_no;
    }
    wait_job(job);
    trace("encode_end");
    if (tracing) {
        std::lock_guard<std::mutex> g(trace_mu);
        std::sort(trace_ev.begin(), trace_ev.end());
        const double t0 = trace_ev.empty() ? 0.0 : trace_ev.front().first;
        for (auto & ev : trace_ev) fprintf(stderr, "[bert trace] %9.1f us  %s\n", ev.first - t0, ev.second.c_str());
    }
    return job.ok;
}

}  // namespace b200
