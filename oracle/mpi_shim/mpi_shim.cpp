/* mpi_shim.cpp — TEST INFRASTRUCTURE ONLY: in-process implementation of the MPI subset declared in mpi.h. */
#include "mpi.h"

#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <exception>
#include <mutex>
#include <thread>
#include <vector>

namespace
{

struct Message
{
    int src, tag;
    std::vector<char> data;
};

struct World
{
    int n = 1;
    std::mutex m;
    std::condition_variable cv;
    std::vector<std::deque<Message>> box; // per destination
    int arrived              = 0;          // barrier
    unsigned long generation = 0;
    std::vector<std::vector<char>> stage; // collective staging
};

World g_world;
thread_local int t_rank = 0;

size_t typeSize(MPI_Datatype t)
{
    switch (t)
    {
        case MPI_CHAR:
        case MPI_UNSIGNED_CHAR: return 1;
        case MPI_SHORT:
        case MPI_UNSIGNED_SHORT: return 2;
        case MPI_INT:
        case MPI_UNSIGNED:
        case MPI_FLOAT: return 4;
        default: return 8;
    }
}

void barrier()
{
    std::unique_lock<std::mutex> lk(g_world.m);
    unsigned long gen = g_world.generation;
    if (++g_world.arrived == g_world.n)
    {
        g_world.arrived = 0;
        ++g_world.generation;
        g_world.cv.notify_all();
    }
    else g_world.cv.wait(lk, [gen] { return g_world.generation != gen; });
}

//! oldest message for me matching (source, tag); call with the lock held.  Returns its position or -1.
int findMessage(int source, int tag)
{
    auto& q = g_world.box[t_rank];
    for (size_t k = 0; k < q.size(); ++k)
        if ((source == MPI_ANY_SOURCE || q[k].src == source) && (tag == MPI_ANY_TAG || q[k].tag == tag)) return int(k);
    return -1;
}

template<class T>
void reduceTyped(T* acc, const T* in, int count, MPI_Op op)
{
    for (int i = 0; i < count; ++i)
    {
        if (op == MPI_SUM) acc[i] = acc[i] + in[i];
        else if (op == MPI_MIN) acc[i] = in[i] < acc[i] ? in[i] : acc[i];
        else acc[i] = in[i] > acc[i] ? in[i] : acc[i];
    }
}

void reduceInto(void* acc, const void* in, int count, MPI_Datatype t, MPI_Op op)
{
    switch (t)
    {
        case MPI_CHAR: reduceTyped((char*)acc, (const char*)in, count, op); break;
        case MPI_UNSIGNED_CHAR: reduceTyped((unsigned char*)acc, (const unsigned char*)in, count, op); break;
        case MPI_SHORT: reduceTyped((short*)acc, (const short*)in, count, op); break;
        case MPI_UNSIGNED_SHORT: reduceTyped((unsigned short*)acc, (const unsigned short*)in, count, op); break;
        case MPI_INT: reduceTyped((int*)acc, (const int*)in, count, op); break;
        case MPI_UNSIGNED: reduceTyped((unsigned*)acc, (const unsigned*)in, count, op); break;
        case MPI_LONG: reduceTyped((long*)acc, (const long*)in, count, op); break;
        case MPI_UNSIGNED_LONG: reduceTyped((unsigned long*)acc, (const unsigned long*)in, count, op); break;
        case MPI_UNSIGNED_LONG_LONG:
            reduceTyped((unsigned long long*)acc, (const unsigned long long*)in, count, op);
            break;
        case MPI_FLOAT: reduceTyped((float*)acc, (const float*)in, count, op); break;
        case MPI_DOUBLE: reduceTyped((double*)acc, (const double*)in, count, op); break;
        default: break;
    }
}

} // namespace

extern "C" {

int MPI_Comm_size(MPI_Comm, int* size)
{
    *size = g_world.n;
    return MPI_SUCCESS;
}
int MPI_Comm_rank(MPI_Comm, int* rank)
{
    *rank = t_rank;
    return MPI_SUCCESS;
}
int MPI_Barrier(MPI_Comm)
{
    barrier();
    return MPI_SUCCESS;
}

int MPI_Isend(const void* buf, int count, MPI_Datatype type, int dest, int tag, MPI_Comm, MPI_Request* req)
{
    Message msg;
    msg.src = t_rank, msg.tag = tag;
    size_t bytes = size_t(count) * typeSize(type);
    msg.data.assign((const char*)buf, (const char*)buf + bytes);
    {
        std::lock_guard<std::mutex> lk(g_world.m);
        g_world.box[dest].push_back(std::move(msg));
    }
    g_world.cv.notify_all();
    if (req) *req = MPI_Request{nullptr, 0, 0, 0, 0};
    return MPI_SUCCESS;
}

int MPI_Probe(int source, int tag, MPI_Comm, MPI_Status* status)
{
    std::unique_lock<std::mutex> lk(g_world.m);
    int k = -1;
    g_world.cv.wait(lk, [&] { return (k = findMessage(source, tag)) >= 0; });
    if (status)
    {
        const Message& msg = g_world.box[t_rank][k];
        status->MPI_SOURCE = msg.src;
        status->MPI_TAG    = msg.tag;
        status->MPI_ERROR  = 0;
        status->shim_bytes = msg.data.size();
    }
    return MPI_SUCCESS;
}

int MPI_Recv(void* buf, int count, MPI_Datatype type, int source, int tag, MPI_Comm, MPI_Status* status)
{
    std::unique_lock<std::mutex> lk(g_world.m);
    int k = -1;
    g_world.cv.wait(lk, [&] { return (k = findMessage(source, tag)) >= 0; });
    auto& q     = g_world.box[t_rank];
    Message msg = std::move(q[k]);
    q.erase(q.begin() + k);
    lk.unlock();
    size_t cap = size_t(count) * typeSize(type);
    if (msg.data.size() > cap)
    {
        std::fprintf(stderr, "mpi_shim: message of %zu bytes truncated to %zu\n", msg.data.size(), cap);
        std::terminate();
    }
    if (!msg.data.empty()) std::memcpy(buf, msg.data.data(), msg.data.size());
    if (status)
    {
        status->MPI_SOURCE = msg.src;
        status->MPI_TAG    = msg.tag;
        status->MPI_ERROR  = 0;
        status->shim_bytes = msg.data.size();
    }
    return MPI_SUCCESS;
}

int MPI_Irecv(void* buf, int count, MPI_Datatype type, int source, int tag, MPI_Comm, MPI_Request* req)
{
    // a message that is already here (the usual Probe -> Irecv sequence, focus/exchange_focus.hpp:82-91) is taken at
    // once, so that the next Probe(MPI_ANY_SOURCE) sees the following one; otherwise MPI_Waitall completes it
    bool here;
    {
        std::lock_guard<std::mutex> lk(g_world.m);
        here = findMessage(source, tag) >= 0;
    }
    if (here)
    {
        MPI_Recv(buf, count, type, source, tag, MPI_COMM_WORLD, MPI_STATUS_IGNORE);
        *req = MPI_Request{nullptr, 0, 0, 0, 0};
    }
    else *req = MPI_Request{buf, size_t(count) * typeSize(type), source, tag, 1};
    return MPI_SUCCESS;
}

int MPI_Get_count(const MPI_Status* status, MPI_Datatype type, int* count)
{
    *count = int(status->shim_bytes / typeSize(type));
    return MPI_SUCCESS;
}

int MPI_Waitall(int count, MPI_Request* reqs, MPI_Status* statuses)
{
    for (int i = 0; i < count; ++i)
    {
        if (reqs[i].pending)
        {
            MPI_Status st;
            MPI_Recv(reqs[i].buf, int(reqs[i].bytes), MPI_CHAR, reqs[i].src, reqs[i].tag, MPI_COMM_WORLD, &st);
            reqs[i].pending = 0;
            if (statuses) statuses[i] = st;
        }
        else if (statuses) statuses[i] = MPI_Status{t_rank, 0, 0, 0};
    }
    return MPI_SUCCESS;
}

int MPI_Allreduce(const void* src, void* dest, int count, MPI_Datatype type, MPI_Op op, MPI_Comm)
{
    size_t bytes   = size_t(count) * typeSize(type);
    const void* in = (src == MPI_IN_PLACE) ? dest : src;
    g_world.stage[t_rank].assign((const char*)in, (const char*)in + bytes);
    barrier();
    if (bytes) std::memcpy(dest, g_world.stage[0].data(), bytes);
    for (int r = 1; r < g_world.n; ++r)
        reduceInto(dest, g_world.stage[r].data(), count, type, op);
    barrier();
    return MPI_SUCCESS;
}

int MPI_Allgatherv(const void* src, int sendCount, MPI_Datatype sendType, void* dest, const int* counts,
                   const int* displs, MPI_Datatype recvType, MPI_Comm)
{
    size_t es = typeSize(recvType);
    if (src == MPI_IN_PLACE)
    {
        const char* mine = (const char*)dest + size_t(displs[t_rank]) * es;
        g_world.stage[t_rank].assign(mine, mine + size_t(counts[t_rank]) * es);
    }
    else g_world.stage[t_rank].assign((const char*)src, (const char*)src + size_t(sendCount) * typeSize(sendType));
    barrier();
    for (int r = 0; r < g_world.n; ++r)
    {
        size_t b = size_t(counts[r]) * es;
        if (b) std::memcpy((char*)dest + size_t(displs[r]) * es, g_world.stage[r].data(), b);
    }
    barrier();
    return MPI_SUCCESS;
}

int mpi_shim_run(int numRanks, void (*fn)(int, void*), void* arg)
{
    g_world.n = numRanks;
    g_world.box.assign(numRanks, {});
    g_world.stage.assign(numRanks, {});
    g_world.arrived = 0;
    std::vector<std::thread> threads;
    for (int r = 0; r < numRanks; ++r)
        threads.emplace_back(
            [r, fn, arg]
            {
                t_rank = r;
                try
                {
                    fn(r, arg);
                }
                catch (const std::exception& e)
                {
                    std::fprintf(stderr, "mpi_shim: rank %d threw: %s\n", r, e.what());
                    std::terminate(); // the other ranks would wait forever
                }
            });
    for (auto& t : threads)
        t.join();
    return 0;
}

} // extern "C"
