/*
 * comm.hpp — the transport of the domain engine behind one small interface, with two implementations:
 *
 *  NcclComm   one process per GPU, NCCL over NVLink / NVSwitch (the product path): ncclAllReduce, ncclAllGather and one
 *             group of ncclSend / ncclRecv per exchange step;
 *  LocalComm  P ranks = P engine instances driven by P host threads of ONE process on ONE GPU: messages are
 *             device-to-device copies, collectives go through a host barrier.  It exists so that the multi-rank path
 *             (particle exchange, focus-tree peer exchanges, halo exchange) can be checked against the reference's
 *             multi-rank Domain on a single-GPU box; it is host-synchronous and not meant to be fast.
 *
 * Replaces the MPI calls of the reference's multi-rank path [ref: primitives/mpi_wrappers.hpp:112-245,
 * domain/domaindecomp_mpi.hpp:85-174, halos/exchange_halos.hpp:27-93, focus/exchange_focus.hpp:61-433].
 * Every rank must call the same sequence of collectives (as with NCCL / MPI).  All pointers are DEVICE pointers.
 */
#pragma once
#include <cuda_runtime.h>
#include <nccl.h>

#include <condition_variable>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>

#include "../../include/cstone_b200_domain.h"

namespace csb
{

enum class CommType
{
    u32,
    i32,
    u64,
    f32,
    f64
};
enum class CommOp
{
    sum,
    min,
    max
};

inline size_t commTypeSize(CommType t) { return (t == CommType::u64 || t == CommType::f64) ? 8 : 4; }

//! one point-to-point message: `bytes` at device pointer `ptr`, to / from rank `peer`
struct CommMsg
{
    int peer;
    void* ptr;
    size_t bytes;
};

class Comm
{
public:
    virtual ~Comm() = default;
    virtual int rank() const = 0;
    virtual int size() const = 0;
    //! in-place reduction of `count` elements at buf over all ranks
    virtual int allReduce(void* buf, size_t count, CommType t, CommOp op, cudaStream_t s) = 0;
    //! recv[r * bytes .. (r+1) * bytes) = send of rank r
    virtual int allGather(const void* send, void* recv, size_t bytes, cudaStream_t s) = 0;
    /*! one exchange step: all sends and receives of this rank.  Messages between the same pair of ranks are matched in
     *  the order they are listed.  Every rank calls it, also with empty lists. */
    virtual int exchange(const std::vector<CommMsg>& sends, const std::vector<CommMsg>& recvs, cudaStream_t s) = 0;

    // ---- peer memory (pack-and-push halo exchange, domain.cu)
    //! true: all ranks are host threads on one device and every step is host-synchronous (kernels must not wait for
    //! a kernel of another rank, they share one stream)
    virtual bool hostSynchronous() const = 0;
    /*! collective: peers[r] = a device pointer through which THIS rank can write the buffer `mine` of rank r
     *  (peers[rank()] = mine).  All ranks call it with buffers from cudaMalloc. */
    virtual int shareBuffer(void* mine, std::vector<void*>& peers, cudaStream_t s) = 0;
    //! undo shareBuffer (before the buffers are freed)
    virtual void releaseShared(std::vector<void*>& peers) = 0;
    //! all ranks have reached this point (host side); a no-op where kernels synchronise through flags instead
    virtual int hostBarrier(cudaStream_t s) = 0;
};

// ------------------------------------------------------------------------------------------------ NCCL
class NcclComm final : public Comm
{
public:
    NcclComm(ncclComm_t c, int rank, int size)
        : comm_(c)
        , rank_(rank)
        , size_(size)
    {
    }
    int rank() const override { return rank_; }
    int size() const override { return size_; }

    int allReduce(void* buf, size_t count, CommType t, CommOp op, cudaStream_t s) override
    {
        static const ncclDataType_t types[] = {ncclUint32, ncclInt32, ncclUint64, ncclFloat, ncclDouble};
        static const ncclRedOp_t ops[]      = {ncclSum, ncclMin, ncclMax};
        return ncclAllReduce(buf, buf, count, types[int(t)], ops[int(op)], comm_, s) == ncclSuccess ? 0 : CSTONE_E_NCCL;
    }
    int allGather(const void* send, void* recv, size_t bytes, cudaStream_t s) override
    {
        return ncclAllGather(send, recv, bytes, ncclInt8, comm_, s) == ncclSuccess ? 0 : CSTONE_E_NCCL;
    }
    int exchange(const std::vector<CommMsg>& sends, const std::vector<CommMsg>& recvs, cudaStream_t s) override
    {
        // the group is always closed, also after a failed call inside it (a dangling group would hang the peers)
        bool ok = ncclGroupStart() == ncclSuccess;
        for (const CommMsg& m : sends)
            if (m.bytes) ok = (ncclSend(m.ptr, m.bytes, ncclInt8, m.peer, comm_, s) == ncclSuccess) && ok;
        for (const CommMsg& m : recvs)
            if (m.bytes) ok = (ncclRecv(m.ptr, m.bytes, ncclInt8, m.peer, comm_, s) == ncclSuccess) && ok;
        ok = (ncclGroupEnd() == ncclSuccess) && ok;
        return ok ? 0 : CSTONE_E_NCCL;
    }

    bool hostSynchronous() const override { return false; }
    //! CUDA IPC: the ranks are processes on one node, the GPUs reach each other over NVLink / NVSwitch
    int shareBuffer(void* mine, std::vector<void*>& peers, cudaStream_t s) override
    {
        // every rank goes through the all-gather, also one that could not export its buffer (the peers would hang)
        cudaIpcMemHandle_t h;
        std::memset(&h, 0, sizeof(h));
        int status = 0;
        if (cudaIpcGetMemHandle(&h, mine) != cudaSuccess) status = int(cudaGetLastError());
        void* stage = nullptr;
        const size_t hb = sizeof(cudaIpcMemHandle_t);
        if (cudaMalloc(&stage, hb * size_t(size_ + 1)) != cudaSuccess) return int(cudaGetLastError());
        std::vector<cudaIpcMemHandle_t> all(size_);
        cudaError_t e = cudaMemcpyAsync(stage, &h, hb, cudaMemcpyHostToDevice, s);
        int st        = e == cudaSuccess ? allGather(stage, static_cast<char*>(stage) + hb, hb, s) : int(e);
        if (st == 0) e = cudaMemcpyAsync(all.data(), static_cast<char*>(stage) + hb, hb * size_, cudaMemcpyDeviceToHost, s);
        if (st == 0 && e == cudaSuccess) e = cudaStreamSynchronize(s);
        cudaFree(stage);
        if (st) return st;
        if (e != cudaSuccess) return int(e);
        if (status) return status;
        peers.assign(size_, nullptr);
        for (int r = 0; r < size_; ++r)
        {
            if (r == rank_) peers[r] = mine;
            else if (cudaIpcOpenMemHandle(&peers[r], all[r], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess)
            {
                int err = int(cudaGetLastError());
                peers[r] = nullptr;
                releaseShared(peers);
                return err;
            }
        }
        return 0;
    }
    void releaseShared(std::vector<void*>& peers) override
    {
        for (int r = 0; r < int(peers.size()); ++r)
            if (r != rank_ && peers[r]) cudaIpcCloseMemHandle(peers[r]);
        peers.clear();
    }
    int hostBarrier(cudaStream_t) override { return 0; }

private:
    ncclComm_t comm_;
    int rank_, size_;
};

// ------------------------------------------------------------------------------------------------ in-process
//! shared state of the P thread-ranks of one process (create once, hand to every engine)
struct LocalWorld
{
    explicit LocalWorld(int n)
        : size(n)
        , ptrs(n, nullptr)
        , sends(n)
    {
    }
    int size;
    std::mutex m;
    std::condition_variable cv;
    int arrived              = 0;
    unsigned long generation = 0;
    std::vector<const void*> ptrs;           // per rank: buffer published for the running collective
    std::vector<std::vector<CommMsg>> sends; // per source rank: messages of the running exchange

    void barrier()
    {
        std::unique_lock<std::mutex> lk(m);
        unsigned long gen = generation;
        if (++arrived == size)
        {
            arrived = 0;
            ++generation;
            cv.notify_all();
        }
        else cv.wait(lk, [this, gen] { return generation != gen; });
    }
};

class LocalComm final : public Comm
{
public:
    LocalComm(LocalWorld* w, int rank)
        : w_(w)
        , rank_(rank)
    {
    }
    int rank() const override { return rank_; }
    int size() const override { return w_->size; }

    int allReduce(void* buf, size_t count, CommType t, CommOp op, cudaStream_t s) override
    {
        const size_t bytes = count * commTypeSize(t);
        if (cudaStreamSynchronize(s) != cudaSuccess) return int(cudaGetLastError());
        w_->ptrs[rank_] = buf;
        w_->barrier();
        // every rank reduces all contributions in rank order (same result everywhere), then writes its own buffer
        std::vector<char> acc(bytes), in(bytes);
        cudaError_t e = cudaSuccess;
        for (int r = 0; r < w_->size && e == cudaSuccess; ++r)
        {
            e = cudaMemcpy(r == 0 ? acc.data() : in.data(), w_->ptrs[r], bytes, cudaMemcpyDeviceToHost);
            if (r > 0 && e == cudaSuccess) reduce(acc.data(), in.data(), count, t, op);
        }
        w_->barrier(); // all have read before anyone overwrites
        if (e == cudaSuccess) e = cudaMemcpy(buf, acc.data(), bytes, cudaMemcpyHostToDevice);
        return int(e);
    }

    int allGather(const void* send, void* recv, size_t bytes, cudaStream_t s) override
    {
        if (cudaStreamSynchronize(s) != cudaSuccess) return int(cudaGetLastError());
        w_->ptrs[rank_] = send;
        w_->barrier();
        cudaError_t e = cudaSuccess;
        for (int r = 0; r < w_->size && e == cudaSuccess; ++r)
            e = cudaMemcpyAsync(static_cast<char*>(recv) + size_t(r) * bytes, w_->ptrs[r], bytes, cudaMemcpyDeviceToDevice, s);
        if (e == cudaSuccess) e = cudaStreamSynchronize(s);
        w_->barrier(); // senders may reuse their buffers now
        return int(e);
    }

    int exchange(const std::vector<CommMsg>& sends, const std::vector<CommMsg>& recvs, cudaStream_t s) override
    {
        if (cudaStreamSynchronize(s) != cudaSuccess) return int(cudaGetLastError());
        w_->sends[rank_] = sends;
        w_->barrier();
        cudaError_t e = cudaSuccess;
        std::vector<size_t> next(w_->size, 0); // per source: position of the next unmatched message addressed to me
        int status = 0;
        for (const CommMsg& rm : recvs)
        {
            const std::vector<CommMsg>& from = w_->sends[rm.peer];
            size_t& k                        = next[rm.peer];
            while (k < from.size() && from[k].peer != rank_)
                ++k;
            if (k == from.size() || from[k].bytes != rm.bytes)
            {
                status = CSTONE_E_SIZE_MISMATCH; // unmatched receive: a protocol error of the engine
                break;
            }
            if (rm.bytes && e == cudaSuccess)
                e = cudaMemcpyAsync(rm.ptr, from[k].ptr, rm.bytes, cudaMemcpyDeviceToDevice, s);
            ++k;
        }
        if (e == cudaSuccess) e = cudaStreamSynchronize(s);
        w_->barrier();
        return status ? status : int(e);
    }

    bool hostSynchronous() const override { return true; }
    //! one process, one device: the peers' pointers are usable as they are
    int shareBuffer(void* mine, std::vector<void*>& peers, cudaStream_t s) override
    {
        if (cudaStreamSynchronize(s) != cudaSuccess) return int(cudaGetLastError());
        w_->ptrs[rank_] = mine;
        w_->barrier();
        peers.assign(w_->size, nullptr);
        for (int r = 0; r < w_->size; ++r)
            peers[r] = const_cast<void*>(w_->ptrs[r]);
        w_->barrier();
        // test hook: the last rank behaves as if the peers' memory could not be mapped (the engine must then fall back
        // to send/recv on ALL ranks)
        if (std::getenv("CSB_TEST_FAIL_SHARE") && rank_ == w_->size - 1)
        {
            peers.clear();
            return int(cudaErrorPeerAccessUnsupported);
        }
        return 0;
    }
    void releaseShared(std::vector<void*>& peers) override { peers.clear(); }
    int hostBarrier(cudaStream_t s) override
    {
        cudaError_t e = cudaStreamSynchronize(s);
        w_->barrier();
        return int(e);
    }

private:
    template<class T>
    static void reduceTyped(T* acc, const T* in, size_t n, CommOp op)
    {
        for (size_t i = 0; i < n; ++i)
        {
            if (op == CommOp::sum) acc[i] = T(acc[i] + in[i]);
            else if (op == CommOp::min) acc[i] = in[i] < acc[i] ? in[i] : acc[i];
            else acc[i] = in[i] > acc[i] ? in[i] : acc[i];
        }
    }
    static void reduce(void* acc, const void* in, size_t n, CommType t, CommOp op)
    {
        switch (t)
        {
            case CommType::u32: reduceTyped((uint32_t*)acc, (const uint32_t*)in, n, op); break;
            case CommType::i32: reduceTyped((int32_t*)acc, (const int32_t*)in, n, op); break;
            case CommType::u64: reduceTyped((uint64_t*)acc, (const uint64_t*)in, n, op); break;
            case CommType::f32: reduceTyped((float*)acc, (const float*)in, n, op); break;
            case CommType::f64: reduceTyped((double*)acc, (const double*)in, n, op); break;
        }
    }
    LocalWorld* w_;
    int rank_;
};

} // namespace csb
