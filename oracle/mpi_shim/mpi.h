/*
 * mpi.h — TEST INFRASTRUCTURE ONLY (oracle build): a single-process subset of MPI in which a rank is a thread.
 *
 * The reference's multi-rank CPU path (Domain<KeyType,T,CpuTag>, domain/domain.hpp:196-243) includes <mpi.h>; no MPI
 * is installed in this image.  This header declares exactly the calls that path makes, and
 * mpi_shim.cpp implements them over in-process mailboxes so that P thread-ranks run the UNMODIFIED reference
 * headers inside one process: Isend = eager copy into the destination's mailbox, Probe/Recv = blocking dequeue
 * (FIFO per source and tag; MPI_ANY_SOURCE takes the oldest message with the tag), Irecv takes a message that is
 * already there at once and is otherwise completed by Waitall, collectives go through a staging area between two
 * barriers and reduce in rank order.  MPI_COMM_WORLD is the only communicator.  Nothing in the product links this.
 */
#pragma once
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Op;

typedef struct MPI_Status
{
    int MPI_SOURCE;
    int MPI_TAG;
    int MPI_ERROR;
    size_t shim_bytes;
} MPI_Status;

typedef struct MPI_Request
{
    void* buf;
    size_t bytes;
    int src;
    int tag;
    int pending; /* 1 = receive still to be matched */
} MPI_Request;

#define MPI_COMM_WORLD 0
#define MPI_SUCCESS 0
#define MPI_IN_PLACE ((void*)1)
#define MPI_ANY_SOURCE (-1)
#define MPI_ANY_TAG (-1)
#define MPI_STATUS_IGNORE ((MPI_Status*)0)
#define MPI_STATUSES_IGNORE ((MPI_Status*)0)

enum
{
    MPI_CHAR = 1,
    MPI_UNSIGNED_CHAR,
    MPI_SHORT,
    MPI_UNSIGNED_SHORT,
    MPI_INT,
    MPI_UNSIGNED,
    MPI_LONG,
    MPI_UNSIGNED_LONG,
    MPI_UNSIGNED_LONG_LONG,
    MPI_FLOAT,
    MPI_DOUBLE
};
enum
{
    MPI_SUM = 1,
    MPI_MIN,
    MPI_MAX
};

int MPI_Comm_size(MPI_Comm comm, int* size);
int MPI_Comm_rank(MPI_Comm comm, int* rank);
int MPI_Barrier(MPI_Comm comm);
int MPI_Isend(const void* buf, int count, MPI_Datatype type, int dest, int tag, MPI_Comm comm, MPI_Request* req);
int MPI_Irecv(void* buf, int count, MPI_Datatype type, int source, int tag, MPI_Comm comm, MPI_Request* req);
int MPI_Recv(void* buf, int count, MPI_Datatype type, int source, int tag, MPI_Comm comm, MPI_Status* status);
int MPI_Probe(int source, int tag, MPI_Comm comm, MPI_Status* status);
int MPI_Get_count(const MPI_Status* status, MPI_Datatype type, int* count);
int MPI_Waitall(int count, MPI_Request* reqs, MPI_Status* statuses);
int MPI_Allreduce(const void* src, void* dest, int count, MPI_Datatype type, MPI_Op op, MPI_Comm comm);
int MPI_Allgatherv(const void* src, int sendCount, MPI_Datatype sendType, void* dest, const int* counts,
                   const int* displs, MPI_Datatype recvType, MPI_Comm comm);

/* shim control (not MPI): run fn(rank, arg) on numRanks thread-ranks and join them; returns 0 */
int mpi_shim_run(int numRanks, void (*fn)(int rank, void* arg), void* arg);

#ifdef __cplusplus
}
#endif
