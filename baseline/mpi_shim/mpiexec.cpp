// `mpiexec -np P prog [args]` for the MPI stand-in (baseline tooling; see mpi.h).
//  * stdin goes to rank 0 only, as with a real mpiexec (the reference reads stdin on rank 0, main.c:68-108);
//  * each rank gets CUDA_VISIBLE_DEVICES = its own GPU when SHIM_GPUS=g0,g1,... is set (the
//    reference always opens device 0, cudaFunctions.cu:209);
//  * OMP_NUM_THREADS defaults to 1 (the reference's parallel read loop scrambles rows otherwise,
//    main.c:96-108);
//  * if any rank dies the others are killed, so a crashed rank cannot hang the job.
#include <fcntl.h>
#include <signal.h>
#include <sys/mman.h>
#include <sys/wait.h>
#include <unistd.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "shim_common.h"

int main(int argc, char** argv) {
    int np = 1, i = 1;
    for (; i < argc; ++i) {
        const std::string a = argv[i];
        if ((a == "-np" || a == "-n") && i + 1 < argc) np = std::atoi(argv[++i]);
        else if (a == "-machinefile" && i + 1 < argc) ++i;  // single host only: ignored
        else if (a == "--map-by" && i + 1 < argc) ++i;
        else break;
    }
    if (i >= argc || np < 1) {
        std::fprintf(stderr, "usage: mpiexec -np P program [args]\n");
        return 2;
    }
    const char* sb = std::getenv("SHIM_BYTES");
    const size_t data_bytes = sb ? std::strtoull(sb, nullptr, 10) : (256ull << 20);
    char path[] = "/tmp/sa_mpishim_XXXXXX";
    const int fd = mkstemp(path);
    if (fd < 0 || ftruncate(fd, (off_t)(SHIM_DATA_OFFSET + data_bytes)) != 0) {
        std::perror("mpiexec: mapping file");
        return 1;
    }
    void* m = mmap(nullptr, SHIM_DATA_OFFSET, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    if (m == MAP_FAILED) {
        std::perror("mpiexec: mmap");
        return 1;
    }
    ShimControl* ctl = static_cast<ShimControl*>(m);
    pthread_barrierattr_t at;
    pthread_barrierattr_init(&at);
    pthread_barrierattr_setpshared(&at, PTHREAD_PROCESS_SHARED);
    pthread_barrier_init(&ctl->barrier, &at, (unsigned)np);
    ctl->world = np;
    ctl->data_bytes = data_bytes;

    std::vector<std::string> gpus;
    if (const char* g = std::getenv("SHIM_GPUS")) {
        std::string s(g), cur;
        for (char c : s) {
            if (c == ',') {
                if (!cur.empty()) gpus.push_back(cur);
                cur.clear();
            } else cur.push_back(c);
        }
        if (!cur.empty()) gpus.push_back(cur);
    }

    std::vector<pid_t> kids;
    for (int r = 0; r < np; ++r) {
        const pid_t pid = fork();
        if (pid == 0) {
            setpgid(0, 0);  // own process group, so a failed job can be torn down with its children
            setenv("SHIM_RANK", std::to_string(r).c_str(), 1);
            setenv("SHIM_SIZE", std::to_string(np).c_str(), 1);
            setenv("SHIM_FILE", path, 1);
            setenv("OMP_NUM_THREADS", "1", 0);
            if (!gpus.empty()) setenv("CUDA_VISIBLE_DEVICES", gpus[(size_t)r % gpus.size()].c_str(), 1);
            if (r != 0) {
                const int nul = open("/dev/null", O_RDONLY);
                dup2(nul, 0);
                close(nul);
            }
            execvp(argv[i], argv + i);
            std::perror("mpiexec: exec");
            _exit(127);
        }
        kids.push_back(pid);
    }
    int rc = 0, left = np;
    while (left > 0) {
        int st = 0;
        const pid_t p = wait(&st);
        if (p < 0) break;
        --left;
        const int code = WIFEXITED(st) ? WEXITSTATUS(st) : 128 + (WIFSIGNALED(st) ? WTERMSIG(st) : 0);
        if (code != 0 && rc == 0) {
            rc = code;
            for (pid_t k : kids)
                if (k != p) {
                    kill(-k, SIGKILL);
                    kill(k, SIGKILL);
                }
        }
    }
    unlink(path);
    return rc;
}
