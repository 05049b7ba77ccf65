// `final --gpus N`: the native multi-process launcher (the role of `mpiexec -np N`, makefile:11).
//
// The launcher reads and packs the problem ONCE (host code only), forks one worker per GPU and then
// serves as the workers' switchboard: it forwards the file descriptors of their CUDA allocations to
// each other (SCM_RIGHTS) and answers their votes / barriers (native_wire.h). It never initialises
// CUDA itself, so the workers inherit no driver state through fork(). Worker 0 prints the result
// lines; a worker that dies or reports an error takes the others down with it (the reference's
// exit(1) leaves its peers blocked in a collective, cudaFunctions.cu:15-33).
#include <dlfcn.h>
#include <poll.h>
#include <signal.h>
#include <sys/socket.h>
#include <sys/wait.h>
#include <unistd.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "native_wire.h"
#include "seqalign.h"

namespace {

double wall_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

// The device list the workers choose from: CUDA_VISIBLE_DEVICES when the caller set one, else 0..N-1.
std::vector<std::string> device_names(int gpus) {
    std::vector<std::string> names;
    if (const char* v = std::getenv("CUDA_VISIBLE_DEVICES")) {
        std::string s(v), cur;
        for (char c : s + ",") {
            if (c == ',') {
                if (!cur.empty()) names.push_back(cur);
                cur.clear();
            } else {
                cur.push_back(c);
            }
        }
    }
    if ((int)names.size() < gpus) {
        names.clear();
        for (int i = 0; i < gpus; ++i) names.push_back(std::to_string(i));
    }
    return names;
}

// SA_NATIVE_SELFTEST=1: a stand-in for the CUDA worker that needs no GPU. It walks through every kind of
// round with pipes in place of CUDA allocation handles (each rank's pipe carries its rank number, so a
// descriptor delivered to the wrong rank or in the wrong order is noticed), and rank 0 "solves" with the
// CPU walk of the packed blob. tests/test_native_launch.py runs it with real processes.
int selftest_worker(sa_native_args* a) {
    int p[2];
    if (pipe(p) != 0) return SA_ERR_ARG;
    const unsigned char tag = (unsigned char)(a->rank + 1);
    if (write(p[1], &tag, 1) != 1) return SA_ERR_ARG;
    close(p[1]);
    sa_wire_msg m = {SA_WIRE_FDS, (uint32_t)a->rank, 0, 1, 0};
    int fds[SA_WIRE_MAX_FDS];
    if (sa_wire_send(a->sock, &m, &p[0]) != 0 || sa_wire_recv(a->sock, &m, fds) != 0 || (int)m.nfds != a->world) return SA_ERR_ARG;
    close(p[0]);
    int64_t ok = 1;
    for (int r = 0; r < a->world; ++r) {
        if (r == a->rank) {  // every rank reads a different rank's pipe exactly once: its right-hand neighbour's
            unsigned char got = 0;
            const int src = (a->rank + 1) % a->world;
            if (read(fds[src], &got, 1) != 1 || got != (unsigned char)(src + 1)) ok = 0;
        }
    }
    for (int r = 0; r < a->world; ++r) close(fds[r]);
    int rp[2] = {-1, -1};
    if (a->rank == 0) {
        if (pipe(rp) != 0) return SA_ERR_ARG;
        const unsigned char all[SA_WIRE_MAX_FDS] = {77, 77, 77, 77, 77, 77, 77, 77, 77, 77, 77, 77, 77, 77, 77, 77};
        if (write(rp[1], all, (size_t)a->world) != a->world) return SA_ERR_ARG;
        close(rp[1]);
    }
    m = {SA_WIRE_ROOT_FD, (uint32_t)a->rank, 0, a->rank == 0 ? 1u : 0u, 0};
    if (sa_wire_send(a->sock, &m, &rp[0]) != 0 || sa_wire_recv(a->sock, &m, fds) != 0 || m.nfds != 1) return SA_ERR_ARG;
    if (a->rank == 0) close(rp[0]);
    unsigned char got = 0;
    if (read(fds[0], &got, 1) != 1 || got != 77) ok = 0;
    close(fds[0]);
    m = {SA_WIRE_MIN, (uint32_t)a->rank, ok ? 100 + a->rank : 0, 0, 0};
    if (sa_wire_send(a->sock, &m, nullptr) != 0 || sa_wire_recv(a->sock, &m, fds) != 0) return SA_ERR_ARG;
    if (m.value != 100) {
        sa_set_last_error("self-test: a descriptor or a vote went to the wrong place");
        return SA_ERR_ARG;
    }
    if (std::getenv("SA_NATIVE_SELFTEST_DIE") && a->rank == a->world - 1) _exit(9);  // a worker that vanishes mid-run
    if (a->rank == 0) return sa_blob_execute_cpu(a->blob, a->header->total_bytes, a->results);
    return SA_OK;
}

int worker_main(int rank, int gpus, int sock, const std::string& root, const void* blob, size_t blob_bytes, int repeat,
                const std::string& timing_path, double t_start) {
    // SA_NATIVE_VISIBLE=own: this worker sees only its own GPU (shorter driver start-up)
    int device = rank;
    if (const char* v = std::getenv("SA_NATIVE_VISIBLE")) {
        if (std::strcmp(v, "own") == 0) {
            setenv("CUDA_VISIBLE_DEVICES", device_names(gpus)[rank].c_str(), 1);
            device = 0;
        }
    }
    const bool selftest = std::getenv("SA_NATIVE_SELFTEST") != nullptr;
    const std::string lib = root + "/mpi_openmp_cuda_b200/_lib/libseqalign_cuda.so";
    void* h = selftest ? dlopen(nullptr, RTLD_NOW) : dlopen(lib.c_str(), RTLD_NOW | RTLD_LOCAL);
    typedef int (*worker_fn)(sa_native_args*);
    typedef const char* (*err_fn)(void);
    worker_fn worker = selftest ? &selftest_worker : h ? reinterpret_cast<worker_fn>(dlsym(h, "sa_native_worker")) : nullptr;
    sa_wire_msg done = {SA_WIRE_DONE, (uint32_t)rank, 1, 0, 0};
    if (!worker) {
        std::fprintf(stderr, "final: cannot load %s (%s)\n", lib.c_str(), dlerror());
        sa_wire_send(sock, &done, nullptr);
        return 1;
    }
    const sa_blob_header* hdr = static_cast<const sa_blob_header*>(blob);
    const uint32_t nseq = hdr->nseq;
    std::vector<int32_t> res((size_t)nseq * 4 + 4);
    sa_native_args a;
    std::memset(&a, 0, sizeof(a));
    a.rank = rank;
    a.world = gpus;
    a.sock = sock;
    a.device = device;
    a.blob = rank == 0 ? blob : nullptr;
    a.header = rank == 0 ? hdr : nullptr;
    a.max_seq = nseq ? nseq : 1;
    a.max_blob_bytes = blob_bytes;
    a.results = res.data();
    a.timeout_ms = 0;
    if (const char* v = std::getenv("SA_FABRIC_TIMEOUT_MS")) a.timeout_ms = (uint32_t)std::atoi(v);
    const char* mc = std::getenv("SA_FABRIC_MULTICAST");
    a.use_multicast = (mc && std::strcmp(mc, "0") == 0) ? 0 : 1;
    a.repeat = repeat;
    const int rc = worker(&a);
    if (rc != SA_OK) {
        err_fn err = selftest ? &sa_last_error : reinterpret_cast<err_fn>(dlsym(h, "sa_last_error"));
        std::fprintf(stderr, "final: GPU worker failed: %s\n", err ? err() : "?");
        sa_wire_send(sock, &done, nullptr);
        return 1;
    }
    if (rank == 0) {
        std::vector<int32_t> res3((size_t)nseq * 3 + 3);
        for (uint32_t i = 0; i < nseq; ++i)
            for (int c = 0; c < 3; ++c) res3[(size_t)i * 3 + c] = res[(size_t)i * 4 + c];
        const size_t need = sa_format(res3.data(), (int32_t)nseq, nullptr, 0);
        std::string out(need, '\0');
        sa_format(res3.data(), (int32_t)nseq, out.data(), need);
        fwrite(out.data(), 1, out.size(), stdout);
        fflush(stdout);
        if (!timing_path.empty()) {
            if (FILE* f = std::fopen(timing_path.c_str(), "w")) {
                std::fprintf(f,
                             "{\"engine\": \"gpu-fabric-native\", \"gpus\": %d, \"nseq\": %u, \"multicast\": %d, \"kernel_ms\": %.6f, "
                             "\"cuda_up_s\": %.6f, \"buffers_mapped_s\": %.6f, \"ready_to_launch_s\": %.6f, \"results_on_host_s\": %.6f, "
                             "\"since_process_start_s\": %.6f}\n",
                             gpus, nseq, a.used_multicast, a.kernel_ms, a.times[0], a.times[1], a.times[2], a.times[3], wall_s() - t_start);
                std::fclose(f);
            }
        }
    }
    done.value = 0;
    sa_wire_send(sock, &done, nullptr);
    return 0;
}

}  // namespace

namespace {

// One round: a message of the same type from every worker, then the answer to every worker.
// Returns 1 when the round was the workers' last (DONE), 0 to continue, -1 on failure.
int serve_round(const std::vector<int>& socks, int timeout_ms, std::string& why) {
    const int world = (int)socks.size();
    std::vector<sa_wire_msg> msgs(world);
    std::vector<int> fds_of(world, -1);
    auto close_all = [&]() {
        for (int& fd : fds_of)
            if (fd >= 0) close(fd), fd = -1;
    };
    for (int r = 0; r < world; ++r) {
        struct pollfd p = {socks[r], POLLIN, 0};
        const int pr = poll(&p, 1, timeout_ms);
        int fds[SA_WIRE_MAX_FDS];
        if (pr <= 0 || sa_wire_recv(socks[r], &msgs[r], fds) != 0) {
            why = "worker " + std::to_string(r) + (pr == 0 ? " did not answer in time" : " went away");
            close_all();
            return -1;
        }
        for (uint32_t i = 1; i < msgs[r].nfds; ++i) close(fds[i]);  // a worker sends at most one descriptor per round
        if (msgs[r].nfds >= 1) fds_of[r] = fds[0];
        if (msgs[r].nfds > 1) why = "worker " + std::to_string(r) + " sent a malformed message";
        else if (msgs[r].type == SA_WIRE_DONE && msgs[r].value != 0) why = "worker " + std::to_string(r) + " reported an error";
        else if (msgs[r].type != msgs[0].type) why = "workers are out of step";
        else continue;
        close_all();
        return -1;
    }
    sa_wire_msg reply = msgs[0];
    int rc = 0;
    switch (msgs[0].type) {
        case SA_WIRE_FDS: {
            for (int r = 0; r < world; ++r)
                if (fds_of[r] < 0) {
                    why = "worker " + std::to_string(r) + " sent no descriptor";
                    close_all();
                    return -1;
                }
            reply.nfds = (uint32_t)world;
            for (int r = 0; r < world && rc == 0; ++r) rc = sa_wire_send(socks[r], &reply, fds_of.data());
            break;
        }
        case SA_WIRE_ROOT_FD: {
            reply.nfds = fds_of[0] >= 0 ? 1u : 0u;
            for (int r = 0; r < world && rc == 0; ++r) rc = sa_wire_send(socks[r], &reply, &fds_of[0]);
            break;
        }
        case SA_WIRE_MIN: {
            for (int r = 1; r < world; ++r)
                if (msgs[r].value < reply.value) reply.value = msgs[r].value;
            reply.nfds = 0;
            for (int r = 0; r < world && rc == 0; ++r) rc = sa_wire_send(socks[r], &reply, nullptr);
            break;
        }
        case SA_WIRE_DONE:
            close_all();
            return 1;
        default:
            why = "unknown message type";
            close_all();
            return -1;
    }
    close_all();
    if (rc != 0) {
        why = "a worker went away while being answered";
        return -1;
    }
    return 0;
}

}  // namespace

// Returns the process exit code. `text` is the whole of stdin.
int sa_native_launch(const std::string& text, int gpus, const std::string& root, const std::string& timing_path, int threads) {
    const double t_start = wall_s();
    if (gpus < 1 || gpus > SA_MAX_WORLD) {
        std::fprintf(stderr, "final: --gpus must be 1..%d\n", SA_MAX_WORLD);
        return 2;
    }
    sa_problem* prob = nullptr;
    if (sa_parse(text.data(), text.size(), &prob) != SA_OK) {
        std::fprintf(stderr, "final: %s\n", sa_last_error());
        return 1;
    }
    void* blob = nullptr;
    size_t blob_bytes = 0;
    const int prc = sa_plan_pack(prob, gpus, &blob, &blob_bytes);
    if (prc == SA_ERR_UNSUPPORTED) {  // same fallback as the single-GPU path
        std::fprintf(stderr, "final: input outside the GPU path (%s); using the CPU oracle\n", sa_last_error());
        const int32_t n = sa_problem_nseq(prob);
        std::vector<int32_t> res((size_t)n * 3 + 3);
        if (sa_solve_cpu_fast(prob, res.data(), threads) != SA_OK) return 1;
        const size_t need = sa_format(res.data(), n, nullptr, 0);
        std::string out(need, '\0');
        sa_format(res.data(), n, out.data(), need);
        fwrite(out.data(), 1, out.size(), stdout);
        return 0;
    }
    if (prc != SA_OK) {
        std::fprintf(stderr, "final: %s\n", sa_last_error());
        return 1;
    }
    int repeat = 1;
    if (const char* v = std::getenv("SA_NATIVE_REPEAT")) repeat = std::atoi(v) > 0 ? std::atoi(v) : 1;
    uint32_t timeout_ms = 10000;
    if (const char* v = std::getenv("SA_FABRIC_TIMEOUT_MS")) timeout_ms = (uint32_t)std::atoi(v);

    std::vector<int> parent_end(gpus), child_end(gpus);
    for (int r = 0; r < gpus; ++r) {
        int sv[2];
        if (socketpair(AF_UNIX, SOCK_SEQPACKET, 0, sv) != 0) {
            std::perror("final: socketpair");
            return 1;
        }
        parent_end[r] = sv[0];
        child_end[r] = sv[1];
    }
    fflush(stdout);
    fflush(stderr);
    std::vector<pid_t> pids(gpus, -1);
    for (int r = 0; r < gpus; ++r) {
        const pid_t pid = fork();
        if (pid < 0) {
            std::perror("final: fork");
            for (int q = 0; q < r; ++q) kill(pids[q], SIGKILL);
            return 1;
        }
        if (pid == 0) {
            for (int q = 0; q < gpus; ++q) {
                close(parent_end[q]);
                if (q != r) close(child_end[q]);
            }
            const int rc = worker_main(r, gpus, child_end[r], root, blob, blob_bytes, repeat, timing_path, t_start);
            fflush(stdout);
            _exit(rc);
        }
        pids[r] = pid;
    }
    for (int r = 0; r < gpus; ++r) close(child_end[r]);

    // CUDA start-up of a cold process takes seconds; after that every wait is bounded by the fabric's own timeout
    const int round_timeout_ms = 120000 + 4 * (int)timeout_ms;
    std::string why;
    int state = 0;
    while (state == 0) state = serve_round(parent_end, round_timeout_ms, why);
    int exit_code = 0;
    if (state < 0) {
        std::fprintf(stderr, "final: multi-GPU run aborted: %s\n", why.c_str());
        for (pid_t p : pids) kill(p, SIGKILL);  // exactly the workers forked above
        exit_code = 1;
    }
    for (pid_t p : pids) {
        int st = 0;
        if (waitpid(p, &st, 0) == p && state > 0 && !(WIFEXITED(st) && WEXITSTATUS(st) == 0)) exit_code = 1;
    }
    sa_blob_free(blob);
    sa_problem_free(prob);
    return exit_code;
}
