/* mpi_shim.c -- TEST INFRASTRUCTURE ONLY (see mpi.h).  fork()-based ranks + shared-memory collectives.
 *   MPISHIM_NP      number of ranks (default 1)
 *   MPISHIM_BUF_MB  size of the shared exchange window (default 1024, lazily committed)
 */
#define _GNU_SOURCE
#include "mpi.h"

#include <sched.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/resource.h>
#include <sys/wait.h>
#include <time.h>
#include <unistd.h>

struct shm_ctl {
    volatile int arrived;
    volatile int sense;
};

static int g_rank = 0, g_size = 1;
static struct shm_ctl *g_ctl;
static char *g_buf;
static size_t g_buf_bytes;
static int g_local_sense = 0;
static size_t g_type_size[64] = {0, 1, sizeof(int), sizeof(double)};
static int g_ntypes = 4;

static void barrier(void)
{
    if (g_size == 1) return;
    g_local_sense ^= 1;
    if (__sync_add_and_fetch(&g_ctl->arrived, 1) == g_size) {
        g_ctl->arrived = 0;
        __sync_synchronize();
        g_ctl->sense = g_local_sense;
    } else {
        int spins = 0;
        while (g_ctl->sense != g_local_sense) {
            if (++spins > 1000) sched_yield();
        }
    }
    __sync_synchronize();
}

int MPI_Init(int *argc, char ***argv)
{
    (void)argc; (void)argv;
    const char *np = getenv("MPISHIM_NP");
    const char *mb = getenv("MPISHIM_BUF_MB");
    g_size = np ? atoi(np) : 1;
    if (g_size < 1) g_size = 1;
    g_buf_bytes = (size_t)(mb ? atol(mb) : 1024) << 20;
    /* the reference keeps a whole `struct world` on the stack while building its datatype (nbody-par.c:170) */
    struct rlimit rl;
    if (getrlimit(RLIMIT_STACK, &rl) == 0) {
        rl.rlim_cur = rl.rlim_max;
        setrlimit(RLIMIT_STACK, &rl);
    }
    char *m = mmap(NULL, g_buf_bytes + 4096, PROT_READ | PROT_WRITE, MAP_SHARED | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
    if (m == MAP_FAILED) { perror("mpi shim: mmap"); exit(1); }
    g_ctl = (struct shm_ctl *)m;
    g_ctl->arrived = 0;
    g_ctl->sense = 0;
    g_buf = m + 4096;
    fflush(stdout);
    fflush(stderr);
    for (int r = 1; r < g_size; ++r) {
        pid_t p = fork();
        if (p < 0) { perror("mpi shim: fork"); exit(1); }
        if (p == 0) { g_rank = r; break; }
    }
    return MPI_SUCCESS;
}

int MPI_Finalize(void)
{
    barrier();
    fflush(stdout);
    fflush(stderr);
    if (g_rank != 0) _exit(0);
    while (wait(NULL) > 0) {}
    return MPI_SUCCESS;
}

int MPI_Comm_rank(MPI_Comm comm, int *rank) { (void)comm; *rank = g_rank; return MPI_SUCCESS; }
int MPI_Comm_size(MPI_Comm comm, int *size) { (void)comm; *size = g_size; return MPI_SUCCESS; }
int MPI_Get_address(const void *location, MPI_Aint *address) { *address = (MPI_Aint)location; return MPI_SUCCESS; }

int MPI_Type_create_struct(int count, const int blocklengths[], const MPI_Aint displacements[],
                           const MPI_Datatype types[], MPI_Datatype *newtype)
{
    size_t extent = 0;
    for (int k = 0; k < count; ++k) {
        size_t end = (size_t)displacements[k] + (size_t)blocklengths[k] * g_type_size[types[k]];
        if (end > extent) extent = end;
    }
    extent = (extent + 7) & ~(size_t)7;   /* the structs in question are 8-byte aligned */
    if (g_ntypes >= 64) { fprintf(stderr, "mpi shim: too many datatypes\n"); exit(1); }
    g_type_size[g_ntypes] = extent;
    *newtype = g_ntypes++;
    return MPI_SUCCESS;
}

int MPI_Type_commit(MPI_Datatype *type) { (void)type; return MPI_SUCCESS; }
int MPI_Type_free(MPI_Datatype *type) { *type = MPI_DATATYPE_NULL; return MPI_SUCCESS; }
int MPI_Barrier(MPI_Comm comm) { (void)comm; barrier(); return MPI_SUCCESS; }

static void need(size_t bytes)
{
    if (bytes > g_buf_bytes) {
        fprintf(stderr, "mpi shim: exchange of %zu bytes exceeds the window (set MPISHIM_BUF_MB)\n", bytes);
        exit(1);
    }
}

int MPI_Bcast(void *buffer, int count, MPI_Datatype type, int root, MPI_Comm comm)
{
    (void)comm;
    if (g_size == 1) return MPI_SUCCESS;
    size_t bytes = (size_t)count * g_type_size[type];
    need(bytes);
    if (g_rank == root) memcpy(g_buf, buffer, bytes);
    barrier();
    if (g_rank != root) memcpy(buffer, g_buf, bytes);
    barrier();
    return MPI_SUCCESS;
}

int MPI_Allgatherv(const void *sendbuf, int sendcount, MPI_Datatype sendtype, void *recvbuf,
                   const int recvcounts[], const int displs[], MPI_Datatype recvtype, MPI_Comm comm)
{
    (void)comm; (void)sendcount; (void)sendtype;
    if (g_size == 1) return MPI_SUCCESS;
    const size_t ext = g_type_size[recvtype];
    char *rb = (char *)recvbuf;
    const size_t my_off = (size_t)displs[g_rank] * ext, my_len = (size_t)recvcounts[g_rank] * ext;
    need((size_t)(displs[g_size - 1] + recvcounts[g_size - 1]) * ext);
    memcpy(g_buf + my_off, sendbuf == MPI_IN_PLACE ? rb + my_off : (const char *)sendbuf, my_len);
    barrier();
    for (int r = 0; r < g_size; ++r) {
        if (r == g_rank && sendbuf == MPI_IN_PLACE) continue;
        memcpy(rb + (size_t)displs[r] * ext, g_buf + (size_t)displs[r] * ext, (size_t)recvcounts[r] * ext);
    }
    barrier();
    return MPI_SUCCESS;
}

double MPI_Wtime(void)
{
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec + ts.tv_nsec * 1e-9;
}
