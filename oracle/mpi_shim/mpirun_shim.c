/*
 * mpirun_shim -- launcher for the mini-MPI shim (TEST INFRASTRUCTURE, not product code).
 *
 *   mpirun_shim -np N [--arena-mb M] [--timeout S] [--out-prefix P] prog [args...]
 *
 * Stands in for `mpirun -np N prog args` (how the reference launches `speculative`,
 * /root/reference/examples/speculative/scripts/dolphin_tinyllama_8_nodes.sh) on a box that has no MPI.
 * Creates the shared segment, starts N copies of prog with MPISHIM_{SHM,RANK,SIZE} set, and returns
 * the worst exit status.  If one rank dies abnormally, or the timeout expires, the others are killed.
 * With --out-prefix P every rank's stdout goes to P.<rank>.out (stderr to P.<rank>.err).
 */
#define _GNU_SOURCE
#include <errno.h>
#include <fcntl.h>
#include <signal.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <sys/wait.h>
#include <time.h>
#include <unistd.h>

void   mpishim_init_segment(void * base, size_t total, int nranks);
size_t mpishim_header_bytes(void);

int main(int argc, char ** argv) {
    int          np = 1, timeout_s = 0;
    size_t       arena_mb = 1024;
    const char * prefix = NULL;
    int          i = 1;
    for (; i < argc; i++) {
        if (!strcmp(argv[i], "-np") && i + 1 < argc) np = atoi(argv[++i]);
        else if (!strcmp(argv[i], "--arena-mb") && i + 1 < argc) arena_mb = (size_t) atol(argv[++i]);
        else if (!strcmp(argv[i], "--timeout") && i + 1 < argc) timeout_s = atoi(argv[++i]);
        else if (!strcmp(argv[i], "--out-prefix") && i + 1 < argc) prefix = argv[++i];
        else if (!strcmp(argv[i], "--bind-to") && i + 1 < argc) ++i; /* accepted, ignored */
        else break;
    }
    if (i >= argc || np < 1 || np > 32) {
        fprintf(stderr, "usage: %s -np N [--arena-mb M] [--timeout S] [--out-prefix P] prog [args...]\n", argv[0]);
        return 2;
    }
    char path[128];
    snprintf(path, sizeof(path), "/dev/shm/mpishim.%d", (int) getpid());
    int fd = open(path, O_RDWR | O_CREAT | O_TRUNC, 0600);
    if (fd < 0) { perror("open shm"); return 2; }
    size_t total = mpishim_header_bytes() + (arena_mb << 20);
    if (ftruncate(fd, (off_t) total) != 0) { perror("ftruncate"); return 2; }
    void * base = mmap(NULL, total, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    if (base == MAP_FAILED) { perror("mmap"); return 2; }
    close(fd);
    mpishim_init_segment(base, total, np);

    pid_t pids[32];
    for (int r = 0; r < np; r++) {
        pid_t p = fork();
        if (p == 0) {
            char b[32];
            setenv("MPISHIM_SHM", path, 1);
            snprintf(b, sizeof(b), "%d", r);  setenv("MPISHIM_RANK", b, 1);
            snprintf(b, sizeof(b), "%d", np); setenv("MPISHIM_SIZE", b, 1);
            if (prefix) {
                char f[512];
                snprintf(f, sizeof(f), "%s.%d.out", prefix, r);
                int o = open(f, O_WRONLY | O_CREAT | O_TRUNC, 0644);
                if (o >= 0) { dup2(o, 1); close(o); }
                snprintf(f, sizeof(f), "%s.%d.err", prefix, r);
                o = open(f, O_WRONLY | O_CREAT | O_TRUNC, 0644);
                if (o >= 0) { dup2(o, 2); close(o); }
            }
            execvp(argv[i], &argv[i]);
            perror("execvp");
            _exit(127);
        }
        pids[r] = p;
    }
    int    worst = 0, alive = np;
    time_t t0 = time(NULL);
    while (alive > 0) {
        int   st;
        pid_t p = waitpid(-1, &st, WNOHANG);
        if (p > 0) {
            alive--;
            int code = WIFEXITED(st) ? WEXITSTATUS(st) : 128 + WTERMSIG(st);
            for (int r = 0; r < np; r++) if (pids[r] == p) pids[r] = 0;
            if (code != 0) {
                if (code > worst) worst = code;
                for (int r = 0; r < np; r++) if (pids[r] > 0) kill(pids[r], SIGKILL);
            }
        } else {
            if (timeout_s > 0 && time(NULL) - t0 > timeout_s) {
                fprintf(stderr, "mpirun_shim: timeout after %d s, killing ranks\n", timeout_s);
                for (int r = 0; r < np; r++) if (pids[r] > 0) kill(pids[r], SIGKILL);
                worst = 124;
                timeout_s = 0;
            }
            usleep(2000);
        }
    }
    munmap(base, total);
    unlink(path);
    return worst;
}
