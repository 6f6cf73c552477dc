// CPU check of csrc/shm_group.h: `world` processes meet in the segment, run allgather rounds and verify
// every mailbox; with argv[3] = "absent" the last rank never comes and the others must time out cleanly.
#include "../graph_embeddings_b200/csrc/shm_group.h"
#include <cstdlib>
#include <sys/wait.h>
#include <vector>

int main(int argc, char** argv) {
    const int world = argc > 1 ? atoi(argv[1]) : 2, rounds = argc > 2 ? atoi(argv[2]) : 1000;
    const bool absent = argc > 3 && !strcmp(argv[3], "absent");
    unsigned char id[128];
    srand((unsigned)getpid());
    for (auto& b : id) b = (unsigned char)rand();
    std::vector<pid_t> kids;
    int rank = 0;
    for (int r = 1; r < world; r++) {
        pid_t p = fork();
        if (p == 0) { rank = r; kids.clear(); break; }
        kids.push_back(p);
    }
    int rc = 0;
    if (absent && rank == world - 1) return 0;
    ShmGroup g;
    std::string err;
    if (!g.open(id, rank, world, err)) {
        if (absent) { printf("rank %d: timed out as expected: %s\n", rank, err.c_str()); rc = 0; }
        else { fprintf(stderr, "rank %d: %s\n", rank, err.c_str()); rc = 2; }
    } else if (absent) {
        rc = 3;  // must not succeed
    } else {
        struct Msg { int rank, round; double v[4]; };
        std::vector<Msg> all(world);
        for (int k = 0; k < rounds && rc == 0; k++) {
            Msg m{rank, k, {rank * 1.5, (double)k, 0, 0}};
            if (!g.allgather(&m, all.data(), sizeof m, err)) { fprintf(stderr, "rank %d: %s\n", rank, err.c_str()); rc = 4; break; }
            for (int q = 0; q < world; q++)
                if (all[q].rank != q || all[q].round != k || all[q].v[0] != q * 1.5) { fprintf(stderr, "rank %d round %d: bad mailbox %d\n", rank, k, q); rc = 5; }
        }
        g.close();
    }
    if (rank == 0) {
        for (pid_t p : kids) { int st = 0; waitpid(p, &st, 0); if (!WIFEXITED(st) || WEXITSTATUS(st) != 0) rc = rc ? rc : 10 + (WIFEXITED(st) ? WEXITSTATUS(st) : 99); }
        if (rc == 0) printf("ok world=%d rounds=%d%s\n", world, rounds, absent ? " (absent peer)" : "");
    }
    return rc;
}
