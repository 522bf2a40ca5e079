// server_b200 -- the reference's TCP embedding server (examples/server.cpp:35-125) with the part a GPU
// needs added: requests from all connections are batched into one bert_encode_batch call.
//
// Wire protocol, unchanged (examples/server.cpp:104-116, examples/sample_client.py): on connect the
// server sends int32 n_embd; then every message the client sends (one recv of at most 32 KiB of UTF-8,
// no framing) is answered with n_embd float32.  An empty read closes the connection.
//
// What differs from the reference server:
//   * it accepts any number of simultaneous clients (the reference serves one at a time,
//     server.cpp:100-118) -- one thread per connection;
//   * a single batcher thread owns the bert_ctx: it takes every request that is pending (waiting at most
//     BERT_SERVER_BATCH_US, default 200 us, for more to arrive once it has one, up to
//     BERT_SERVER_MAX_BATCH, default 4096) and encodes them in ONE bert_encode_batch call, so
//     single-sentence traffic reaches GPU-efficient batches;
//   * SIGINT / SIGTERM stop it cleanly (bert_free is called).
// Same command line as the reference (bert_params_parse: -m model, --port, -t threads).
#include "bert.h"

#include <arpa/inet.h>
#include <netinet/tcp.h>
#include <sys/socket.h>
#include <unistd.h>

#include <atomic>
#include <chrono>
#include <condition_variable>
#include <csignal>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

namespace {

struct Request {
    std::string text;
    std::vector<float> embedding;
    bool done = false;
    std::mutex m;
    std::condition_variable cv;
};

struct Batcher {
    bert_ctx * ctx;
    int n_threads, n_embd;
    int max_batch = 4096;
    int wait_us = 200;
    std::mutex m;
    std::condition_variable cv;
    std::deque<Request *> pending;
    std::atomic<bool> stop{false};
    std::atomic<uint64_t> n_requests{0}, n_batches{0};

    void submit(Request * r) {
        { std::lock_guard<std::mutex> g(m); pending.push_back(r); }
        cv.notify_one();
    }

    void run() {
        std::vector<Request *> batch;
        std::vector<const char *> texts;
        std::vector<float *> rows;
        while (true) {
            batch.clear();
            {
                std::unique_lock<std::mutex> g(m);
                cv.wait(g, [&] { return stop.load() || !pending.empty(); });
                if (stop.load() && pending.empty()) return;
                // one request is here: give concurrent clients a short window to join the batch
                const auto deadline = std::chrono::steady_clock::now() + std::chrono::microseconds(wait_us);
                size_t seen = 0;
                while ((int) pending.size() < max_batch && pending.size() != seen) {
                    seen = pending.size();
                    if (cv.wait_until(g, deadline, [&] { return pending.size() != seen || stop.load(); })) continue;
                    break;
                }
                while (!pending.empty() && (int) batch.size() < max_batch) {
                    batch.push_back(pending.front());
                    pending.pop_front();
                }
            }
            texts.resize(batch.size());
            rows.resize(batch.size());
            for (size_t i = 0; i < batch.size(); ++i) {
                batch[i]->embedding.assign(n_embd, 0.f);
                texts[i] = batch[i]->text.c_str();
                rows[i] = batch[i]->embedding.data();
            }
            bert_encode_batch(ctx, n_threads, (int32_t) batch.size(), (int32_t) batch.size(), texts.data(), rows.data());
            n_requests += batch.size();
            n_batches += 1;
            for (Request * r : batch) {
                // notify while holding the lock: the Request lives on the connection thread's stack and may be
                // destroyed as soon as the waiter can observe done == true
                std::lock_guard<std::mutex> g(r->m);
                r->done = true;
                r->cv.notify_one();
            }
        }
    }
};

void serve_connection(int fd, Batcher * b) {
    const int one = 1;
    setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
    const int32_t n_embd = b->n_embd;
    if (send(fd, &n_embd, sizeof(n_embd), MSG_NOSIGNAL) != (ssize_t) sizeof(n_embd)) { close(fd); return; }
    std::vector<char> buffer(1 << 15);                     // the reference's message limit (server.cpp:26)
    while (!b->stop.load()) {
        const ssize_t n = recv(fd, buffer.data(), buffer.size(), 0);
        if (n <= 0) break;
        Request r;
        r.text.assign(buffer.data(), (size_t) n);
        b->submit(&r);
        {
            std::unique_lock<std::mutex> g(r.m);
            r.cv.wait(g, [&] { return r.done; });
        }
        const char * p = reinterpret_cast<const char *>(r.embedding.data());
        size_t left = r.embedding.size() * sizeof(float);
        while (left > 0) {
            const ssize_t s = send(fd, p, left, MSG_NOSIGNAL);
            if (s <= 0) { left = 1; break; }
            p += s; left -= (size_t) s;
        }
        if (left) break;
    }
    close(fd);
}

std::atomic<int> g_listen_fd{-1};
void on_signal(int) {
    const int fd = g_listen_fd.exchange(-1);
    if (fd >= 0) shutdown(fd, SHUT_RDWR);                  // wakes accept()
}

}  // namespace

int main(int argc, char ** argv) {
    bert_params params;
    params.model = "models/all-MiniLM-L6-v2/ggml-model-q4_0.bin";
    if (!bert_params_parse(argc, argv, params)) return 1;

    bert_ctx * ctx = bert_load_from_file(params.model);
    if (ctx == nullptr) {
        fprintf(stderr, "%s: failed to load model from '%s'\n", __func__, params.model);
        return 1;
    }

    const int server_fd = socket(AF_INET, SOCK_STREAM, 0);
    if (server_fd < 0) { perror("socket"); return 1; }
    const int one = 1;
    setsockopt(server_fd, SOL_SOCKET, SO_REUSEADDR, &one, sizeof(one));
    sockaddr_in address{};
    address.sin_family = AF_INET;
    address.sin_addr.s_addr = INADDR_ANY;
    address.sin_port = htons((uint16_t) params.port);
    if (bind(server_fd, (sockaddr *) &address, sizeof(address)) < 0) { perror("bind"); return 1; }
    if (listen(server_fd, 256) < 0) { perror("listen"); return 1; }
    socklen_t alen = sizeof(address);
    getsockname(server_fd, (sockaddr *) &address, &alen);   // --port 0: report the port the kernel chose

    Batcher batcher;
    batcher.ctx = ctx;
    batcher.n_threads = params.n_threads;
    batcher.n_embd = bert_n_embd(ctx);
    if (const char * e = getenv("BERT_SERVER_BATCH_US")) batcher.wait_us = atoi(e);
    if (const char * e = getenv("BERT_SERVER_MAX_BATCH")) batcher.max_batch = atoi(e) > 0 ? atoi(e) : 1;
    std::thread batch_thread([&] { batcher.run(); });

    g_listen_fd = server_fd;
    struct sigaction sa{};
    sa.sa_handler = on_signal;
    sigaction(SIGINT, &sa, nullptr);
    sigaction(SIGTERM, &sa, nullptr);

    printf("Server running on port %d with %d threads\n", (int) ntohs(address.sin_port), params.n_threads);
    fflush(stdout);

    while (true) {
        const int fd = accept(server_fd, nullptr, nullptr);
        if (fd < 0) break;                                  // listening socket shut down by the signal handler
        std::thread(serve_connection, fd, &batcher).detach();   // a connection thread owns nothing the server must reap
    }
    batcher.stop = true;
    batcher.cv.notify_all();
    batch_thread.join();
    printf("served %llu requests in %llu batches\n", (unsigned long long) batcher.n_requests.load(),
           (unsigned long long) batcher.n_batches.load());
    fflush(stdout);
    close(server_fd);
    bert_free(ctx);
    _exit(0);
}
