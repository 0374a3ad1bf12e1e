/* Minimal single-node MPI stand-in used ONLY to run the unmodified reference
 * programs (riemann.cpp, 4main.c) on boxes without an MPI installation.
 * Implements exactly the calls the reference makes:
 *   MPI_Init MPI_Comm_size MPI_Comm_rank MPI_Send MPI_Recv MPI_Reduce
 *   MPI_Barrier MPI_Bcast MPI_Get_processor_name MPI_Finalize
 * over one POSIX shared-memory file and process-shared atomics.  Ranks are
 * started by baseline/mpi_shim/mpirun (env: BQ_MPI_RANK, BQ_MPI_SIZE,
 * BQ_MPI_SHM).  Any number measured through it is labelled "MPI shim".
 * Not part of the b200quad product.
 */
#ifndef BQ_MPI_SHIM_H_
#define BQ_MPI_SHIM_H_

#include <fcntl.h>
#include <sched.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Op;
typedef struct { int MPI_SOURCE, MPI_TAG, MPI_ERROR; } MPI_Status;

#define MPI_COMM_WORLD 0
#define MPI_CHAR 1
#define MPI_INT 4
#define MPI_UNSIGNED 5
#define MPI_DOUBLE 8
#define MPI_SUM 1
#define MPI_SUCCESS 0
#define MPI_STATUS_IGNORE ((MPI_Status*)0)
#define MPI_MAX_PROCESSOR_NAME 256

#define BQM_CHUNK (1u << 20)          /* p2p channel buffer */
#define BQM_BCHUNK (8u << 20)         /* broadcast staging buffer (x2) */
#define BQM_MAXP 64
#define BQM_REDUCE_MAX 64

typedef struct {
    volatile uint64_t seq_w, seq_r;   /* chunks written / read */
    volatile uint64_t total;          /* message bytes (valid with its first chunk) */
    volatile uint64_t len;            /* bytes in this chunk */
    char pad[32];
    char buf[BQM_CHUNK];
} bqm_chan;

typedef struct {
    volatile int bar_count;
    volatile int bar_sense;
    char pad0[56];
    volatile uint64_t bc_seq[2];      /* broadcast: generation written into buffer k */
    volatile uint64_t bc_len[2];
    volatile int bc_done[2];          /* readers finished with buffer k */
    char pad1[64];
    double reduce_slot[BQM_MAXP][BQM_REDUCE_MAX];
    char bc_buf[2][BQM_BCHUNK];
} bqm_head;

static bqm_head* bqm_h = 0;
static bqm_chan* bqm_c = 0;           /* [src * size + dst] */
static int bqm_rank = 0, bqm_size = 1, bqm_local_sense = 0;
static uint64_t bqm_bc_gen = 0;

static inline size_t bqm_dtsize(MPI_Datatype t) { return t == MPI_DOUBLE ? 8 : (t == MPI_CHAR ? 1 : 4); }
static inline void bqm_relax(void) {
#if defined(__x86_64__)
    __builtin_ia32_pause();
#endif
}
static inline void bqm_wait_yield(int* spins) {
    if (++*spins > 2000) { sched_yield(); *spins = 0; } else bqm_relax();
}

static inline int MPI_Init(int* argc, char*** argv) {
    (void)argc; (void)argv;
    const char* r = getenv("BQ_MPI_RANK");
    const char* s = getenv("BQ_MPI_SIZE");
    const char* f = getenv("BQ_MPI_SHM");
    bqm_rank = r ? atoi(r) : 0;
    bqm_size = s ? atoi(s) : 1;
    if (bqm_size > BQM_MAXP) { fprintf(stderr, "mpi shim: at most %d ranks\n", BQM_MAXP); exit(1); }
    size_t bytes = sizeof(bqm_head) + sizeof(bqm_chan) * (size_t)bqm_size * (size_t)bqm_size;
    void* p;
    if (f) {
        int fd = open(f, O_RDWR | O_CREAT, 0600);
        if (fd < 0 || ftruncate(fd, (off_t)bytes) != 0) { perror("mpi shim: shm"); exit(1); }
        p = mmap(0, bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
        close(fd);
    } else {
        p = mmap(0, bytes, PROT_READ | PROT_WRITE, MAP_SHARED | MAP_ANONYMOUS, -1, 0);
    }
    if (p == MAP_FAILED) { perror("mpi shim: mmap"); exit(1); }
    bqm_h = (bqm_head*)p;
    bqm_c = (bqm_chan*)((char*)p + sizeof(bqm_head));
    return MPI_SUCCESS;
}

static inline int MPI_Comm_size(MPI_Comm c, int* n) { (void)c; *n = bqm_size; return MPI_SUCCESS; }
static inline int MPI_Comm_rank(MPI_Comm c, int* r) { (void)c; *r = bqm_rank; return MPI_SUCCESS; }

static inline int MPI_Barrier(MPI_Comm c) {
    (void)c;
    if (bqm_size == 1) return MPI_SUCCESS;
    int spins = 0;
    bqm_local_sense = !bqm_local_sense;
    if (__atomic_add_fetch(&bqm_h->bar_count, 1, __ATOMIC_ACQ_REL) == bqm_size) {
        __atomic_store_n(&bqm_h->bar_count, 0, __ATOMIC_RELAXED);
        __atomic_store_n(&bqm_h->bar_sense, bqm_local_sense, __ATOMIC_RELEASE);
    } else {
        while (__atomic_load_n(&bqm_h->bar_sense, __ATOMIC_ACQUIRE) != bqm_local_sense) bqm_wait_yield(&spins);
    }
    return MPI_SUCCESS;
}

static inline int MPI_Send(const void* buf, int count, MPI_Datatype t, int dest, int tag, MPI_Comm c) {
    (void)tag; (void)c;
    bqm_chan* ch = &bqm_c[(size_t)bqm_rank * bqm_size + dest];
    uint64_t total = (uint64_t)count * bqm_dtsize(t), off = 0;
    int spins = 0;
    do {
        uint64_t n = total - off < BQM_CHUNK ? total - off : BQM_CHUNK;
        while (__atomic_load_n(&ch->seq_w, __ATOMIC_RELAXED) != __atomic_load_n(&ch->seq_r, __ATOMIC_ACQUIRE))
            bqm_wait_yield(&spins);
        memcpy(ch->buf, (const char*)buf + off, n);
        ch->total = total;
        ch->len = n;
        __atomic_store_n(&ch->seq_w, ch->seq_w + 1, __ATOMIC_RELEASE);
        off += n;
    } while (off < total);
    return MPI_SUCCESS;
}

static inline int MPI_Recv(void* buf, int count, MPI_Datatype t, int src, int tag, MPI_Comm c, MPI_Status* st) {
    (void)tag; (void)c;
    bqm_chan* ch = &bqm_c[(size_t)src * bqm_size + bqm_rank];
    uint64_t cap = (uint64_t)count * bqm_dtsize(t), off = 0, total = 0;
    int spins = 0, first = 1;
    do {
        while (__atomic_load_n(&ch->seq_w, __ATOMIC_ACQUIRE) == __atomic_load_n(&ch->seq_r, __ATOMIC_RELAXED))
            bqm_wait_yield(&spins);
        if (first) { total = ch->total; first = 0; }
        uint64_t n = ch->len;
        if (off < cap) memcpy((char*)buf + off, ch->buf, off + n <= cap ? n : cap - off);
        off += n;
        __atomic_store_n(&ch->seq_r, ch->seq_r + 1, __ATOMIC_RELEASE);
    } while (off < total);
    if (st) { st->MPI_SOURCE = src; st->MPI_TAG = tag; st->MPI_ERROR = 0; }
    return MPI_SUCCESS;
}

static inline int MPI_Reduce(const void* send, void* recv, int count, MPI_Datatype t, MPI_Op op, int root, MPI_Comm c) {
    (void)op;
    if (t != MPI_DOUBLE || count > BQM_REDUCE_MAX) { fprintf(stderr, "mpi shim: Reduce supports <=%d doubles\n", BQM_REDUCE_MAX); exit(1); }
    memcpy((void*)bqm_h->reduce_slot[bqm_rank], send, sizeof(double) * (size_t)count);
    MPI_Barrier(c);
    if (bqm_rank == root) {
        double* out = (double*)recv;
        for (int i = 0; i < count; ++i) {
            double s = 0.0;
            for (int r = 0; r < bqm_size; ++r) s += bqm_h->reduce_slot[r][i];
            out[i] = s;
        }
    }
    MPI_Barrier(c);
    return MPI_SUCCESS;
}

static inline int MPI_Bcast(void* buf, int count, MPI_Datatype t, int root, MPI_Comm c) {
    (void)c;
    if (bqm_size == 1) return MPI_SUCCESS;
    uint64_t total = (uint64_t)count * bqm_dtsize(t), off = 0;
    int spins = 0;
    while (off < total) {
        uint64_t n = total - off < BQM_BCHUNK ? total - off : BQM_BCHUNK;
        uint64_t gen = ++bqm_bc_gen;          /* same sequence on every rank */
        int k = (int)(gen & 1);
        if (bqm_rank == root) {
            /* buffer k was last used by generation gen-2: wait for its readers */
            if (gen > 2)
                while (__atomic_load_n(&bqm_h->bc_done[k], __ATOMIC_ACQUIRE) != bqm_size - 1) bqm_wait_yield(&spins);
            __atomic_store_n(&bqm_h->bc_done[k], 0, __ATOMIC_RELAXED);
            memcpy(bqm_h->bc_buf[k], (char*)buf + off, n);
            bqm_h->bc_len[k] = n;
            __atomic_store_n(&bqm_h->bc_seq[k], gen, __ATOMIC_RELEASE);
        } else {
            while (__atomic_load_n(&bqm_h->bc_seq[k], __ATOMIC_ACQUIRE) != gen) bqm_wait_yield(&spins);
            memcpy((char*)buf + off, bqm_h->bc_buf[k], n);
            __atomic_add_fetch(&bqm_h->bc_done[k], 1, __ATOMIC_ACQ_REL);
        }
        off += n;
    }
    return MPI_SUCCESS;
}

static inline int MPI_Get_processor_name(char* name, int* len) {
    if (gethostname(name, MPI_MAX_PROCESSOR_NAME - 1) != 0) strcpy(name, "localhost");
    name[MPI_MAX_PROCESSOR_NAME - 1] = 0;
    *len = (int)strlen(name);
    return MPI_SUCCESS;
}

static inline int MPI_Finalize(void) { MPI_Barrier(MPI_COMM_WORLD); return MPI_SUCCESS; }

#ifdef __cplusplus
}
#endif
#endif /* BQ_MPI_SHIM_H_ */
