// Minimal single-host stand-in for the eight MPI functions the reference's
// kdtree_mpi.cpp calls.  Test/baseline tooling only: it lets
// the UNMODIFIED reference translation unit compile and run on a box without
// an MPI toolchain.  MPI_Init forks world-1 children (world = $KD_MPI_NP,
// default 4); collectives go through one anonymous shared mapping guarded by
// a process-shared pthread barrier.
#pragma once
#include <pthread.h>
#include <sys/mman.h>
#include <sys/wait.h>
#include <unistd.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>

typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Op;
#define MPI_COMM_WORLD 0
#define MPI_INT 4
#define MPI_FLOAT 104
#define MPI_MIN 1
#define MPI_SUCCESS 0

namespace kd_mpi_shim {
struct Shared {
    pthread_barrier_t bar;
    unsigned char bcast[4096];
    unsigned char slots[256][4096];
};
inline Shared*& shm() { static Shared* s = nullptr; return s; }
inline int& rank() { static int r = 0; return r; }
inline int& size() { static int s = 1; return s; }
inline int& next_type() { static int t = 1000; return t; }
inline int type_bytes(MPI_Datatype t) {
    if (t == MPI_INT || t == MPI_FLOAT) return 4;
    if (t >= 1000) return (t - 1000) % 100000;  // encoded by MPI_Type_contiguous
    return 4;
}
inline void barrier() { pthread_barrier_wait(&shm()->bar); }
}  // namespace kd_mpi_shim

inline int MPI_Init(int*, char***) {
    using namespace kd_mpi_shim;
    const char* e = getenv("KD_MPI_NP");
    int np = e ? atoi(e) : 4;
    if (np < 1) np = 1;
    if (np > 256) np = 256;
    size() = np;
    void* p = mmap(nullptr, sizeof(Shared), PROT_READ | PROT_WRITE, MAP_SHARED | MAP_ANONYMOUS, -1, 0);
    if (p == MAP_FAILED) { perror("mmap"); exit(2); }
    shm() = (Shared*)p;
    pthread_barrierattr_t a;
    pthread_barrierattr_init(&a);
    pthread_barrierattr_setpshared(&a, PTHREAD_PROCESS_SHARED);
    pthread_barrier_init(&shm()->bar, &a, np);
    fflush(stdout);
    fflush(stderr);
    for (int r = 1; r < np; ++r) {
        pid_t c = fork();
        if (c < 0) { perror("fork"); exit(2); }
        if (c == 0) { rank() = r; break; }
    }
    return MPI_SUCCESS;
}
inline int MPI_Comm_rank(MPI_Comm, int* r) { *r = kd_mpi_shim::rank(); return MPI_SUCCESS; }
inline int MPI_Comm_size(MPI_Comm, int* s) { *s = kd_mpi_shim::size(); return MPI_SUCCESS; }
inline int MPI_Type_contiguous(int count, MPI_Datatype old, MPI_Datatype* out) {
    *out = 1000 + count * kd_mpi_shim::type_bytes(old);
    return MPI_SUCCESS;
}
inline int MPI_Type_commit(MPI_Datatype*) { return MPI_SUCCESS; }
inline int MPI_Bcast(void* buf, int count, MPI_Datatype t, int root, MPI_Comm) {
    using namespace kd_mpi_shim;
    size_t nb = (size_t)count * type_bytes(t);
    if (nb > sizeof(shm()->bcast)) { fprintf(stderr, "shim: bcast too large\n"); exit(2); }
    if (rank() == root) memcpy(shm()->bcast, buf, nb);
    barrier();
    if (rank() != root) memcpy(buf, shm()->bcast, nb);
    barrier();
    return MPI_SUCCESS;
}
inline int MPI_Reduce(const void* send, void* recv, int count, MPI_Datatype t, MPI_Op op, int root, MPI_Comm) {
    using namespace kd_mpi_shim;
    if (t != MPI_FLOAT || op != MPI_MIN || (size_t)count * 4 > sizeof(shm()->slots[0])) {
        fprintf(stderr, "shim: only MPI_Reduce(MPI_FLOAT, MPI_MIN) is provided\n");
        exit(2);
    }
    memcpy(shm()->slots[rank()], send, (size_t)count * 4);
    barrier();
    if (rank() == root) {
        float* out = (float*)recv;
        for (int i = 0; i < count; ++i) {
            float m = ((float*)shm()->slots[0])[i];
            for (int r = 1; r < size(); ++r) {
                float v = ((float*)shm()->slots[r])[i];
                if (v < m) m = v;
            }
            out[i] = m;
        }
    }
    barrier();
    return MPI_SUCCESS;
}
inline int MPI_Finalize() {
    using namespace kd_mpi_shim;
    fflush(stdout);
    fflush(stderr);
    barrier();
    if (rank() != 0) _exit(0);
    for (int r = 1; r < size(); ++r) { int st; wait(&st); }
    return MPI_SUCCESS;
}
