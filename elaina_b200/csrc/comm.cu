// NCCL inside the library: the gradient all-reduce of the replicated guiding network and the final film gather of the
// sharded walker (north_star: "NCCL over NVLink is used only for the gradient allreduce ... and the final grid gather";
// SURVEY 8(e)). The reference is single-device (core/device/context.cpp:23-24), so there is no interface to mirror:
// the entry points are the ones a multi-GPU exec.cu would need -- create a communicator, attach it to an integrator,
// gather the film -- and elaina-exec uses them when it is started by torchrun / mpirun (exec.cpp).
//
// libnccl is resolved at run time (dlopen "libnccl.so.2"): inside a Python process that already loaded torch's bundled
// NCCL the same copy is reused, a standalone elaina-exec picks up the system library. A missing NCCL fails loudly with
// WOST_ERR_UNSUPPORTED at wost_comm_create -- single-GPU use never touches it.
#include "integrator.cuh"
#include "scene.cuh"

#include <dlfcn.h>
#include <fcntl.h>
#include <cctype>
#include <nccl.h>
#include <sys/stat.h>
#include <unistd.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>

namespace wost
{
    void set_last_error(const std::string &s);

    struct NcclApi
    {
        void *handle = nullptr;
        ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
        ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
        ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
        ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
        ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
        ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
        ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
        ncclResult_t (*GroupStart)() = nullptr;
        ncclResult_t (*GroupEnd)() = nullptr;
        const char *(*GetErrorString)(ncclResult_t) = nullptr;
        ncclResult_t (*GetVersion)(int *) = nullptr;
        std::string path;
    };
    static NcclApi *nccl_api()
    {
        static NcclApi api;
        static bool tried = false;
        if (tried)
            return api.handle ? &api : nullptr;
        tried = true;
        const char *cands[] = {getenv("WOST_NCCL_LIB"), "libnccl.so.2", "libnccl.so", "/usr/lib/x86_64-linux-gnu/libnccl.so.2"};
        for (const char *c : cands)
        {
            if (!c || !*c)
                continue;
            if (void *h = dlopen(c, RTLD_NOW | RTLD_GLOBAL))
            {
                api.handle = h;
                api.path = c;
                break;
            }
        }
        if (!api.handle)
            return nullptr;
#define WOST_NCCL_SYM(name) api.name = reinterpret_cast<decltype(api.name)>(dlsym(api.handle, "nccl" #name))
        WOST_NCCL_SYM(GetUniqueId);
        WOST_NCCL_SYM(CommInitRank);
        WOST_NCCL_SYM(CommDestroy);
        WOST_NCCL_SYM(AllReduce);
        WOST_NCCL_SYM(AllGather);
        WOST_NCCL_SYM(Send);
        WOST_NCCL_SYM(Recv);
        WOST_NCCL_SYM(GroupStart);
        WOST_NCCL_SYM(GroupEnd);
        WOST_NCCL_SYM(GetErrorString);
        WOST_NCCL_SYM(GetVersion);
#undef WOST_NCCL_SYM
        if (!api.GetUniqueId || !api.CommInitRank || !api.CommDestroy || !api.AllReduce || !api.AllGather || !api.Send || !api.Recv || !api.GroupStart ||
            !api.GroupEnd || !api.GetErrorString)
        {
            dlclose(api.handle);
            api.handle = nullptr;
            return nullptr;
        }
        return &api;
    }
}

struct wost_comm
{
    ncclComm_t comm = nullptr;
    int rank = 0, world = 1, device = 0;
    wost::DevBuf<float> pack, stage; // film gather: this rank's owned rows / the rows received from one peer
    wost::DevBuf<double> scal;
    std::string id_file;
};

using namespace wost;

#define WOST_NCCL_TRY(expr)                                                                              \
    do                                                                                                   \
    {                                                                                                    \
        ncclResult_t r_ = (expr);                                                                        \
        if (r_ != ncclSuccess)                                                                           \
        {                                                                                                \
            set_last_error(std::string("NCCL: ") + nccl_api()->GetErrorString(r_) + " in " #expr);        \
            return WOST_ERR_CUDA;                                                                        \
        }                                                                                                \
    } while (0)

namespace wost
{
    // rows of the frame a shard owns, in order: tile t = row / stride_rows belongs to shard t % num_shards
    static int owned_row_count(int H, int stride_rows, int num_shards, int shard)
    {
        int n = 0;
        for (int y = 0; y < H; ++y)
            if ((y / stride_rows) % num_shards == shard)
                ++n;
        return n;
    }
    // j-th owned row of `shard` -> frame row
    WOST_HOSTDEV int owned_row(int j, int stride_rows, int num_shards, int shard)
    {
        const int t = j / stride_rows;
        return (t * num_shards + shard) * stride_rows + j % stride_rows;
    }
    // pack (to_packed = 1): packed[j][*] = film[row_j][*]; unpack (0): film[row_j][*] = packed[j][*]; row = W*5 floats
    __global__ void k_film_rows(float *film, float *packed, int row_floats, int nrows, int stride_rows, int num_shards, int shard, int to_packed)
    {
        const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
        if (i >= (long long)nrows * row_floats)
            return;
        const int j = (int)(i / row_floats), k = (int)(i - (long long)j * row_floats);
        const size_t f = (size_t)owned_row(j, stride_rows, num_shards, shard) * row_floats + k;
        if (to_packed)
            packed[i] = film[f];
        else
            film[f] = packed[i];
    }
    static int comm_allreduce_cb(void *user, void *d_buf, uint64_t n_floats, void *stream)
    {
        wost_comm *c = static_cast<wost_comm *>(user);
        NcclApi *api = nccl_api();
        if (!c || !api)
            return 1;
        const ncclResult_t r = api->AllReduce(d_buf, d_buf, (size_t)n_floats, ncclFloat, ncclSum, c->comm, (cudaStream_t)stream);
        if (r != ncclSuccess)
        {
            set_last_error(std::string("NCCL all-reduce of the gradient slab: ") + api->GetErrorString(r));
            return 1;
        }
        return 0;
    }
}

extern "C"
{
    wost_status wost_comm_unique_id(void *id128)
    {
        NcclApi *api = nccl_api();
        if (!id128)
            return WOST_ERR_INVALID;
        if (!api)
        {
            set_last_error("wost_comm: libnccl.so.2 not found (set WOST_NCCL_LIB)");
            return WOST_ERR_UNSUPPORTED;
        }
        static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId size");
        ncclUniqueId id;
        WOST_NCCL_TRY(api->GetUniqueId(&id));
        memcpy(id128, &id, sizeof(id));
        return WOST_OK;
    }

    wost_status wost_comm_create(int rank, int world, const void *id128, wost_comm **out)
    {
        NcclApi *api = nccl_api();
        if (!out || !id128 || world < 1 || rank < 0 || rank >= world)
        {
            set_last_error("wost_comm_create: invalid argument");
            return WOST_ERR_INVALID;
        }
        if (!api)
        {
            set_last_error("wost_comm: libnccl.so.2 not found (set WOST_NCCL_LIB)");
            return WOST_ERR_UNSUPPORTED;
        }
        wost_comm *c = new wost_comm();
        c->rank = rank;
        c->world = world;
        cudaGetDevice(&c->device);
        ncclUniqueId id;
        memcpy(&id, id128, sizeof(id));
        const ncclResult_t r = api->CommInitRank(&c->comm, world, id, rank);
        if (r != ncclSuccess)
        {
            set_last_error(std::string("ncclCommInitRank: ") + api->GetErrorString(r));
            delete c;
            return WOST_ERR_CUDA;
        }
        *out = c;
        return WOST_OK;
    }

    // Launcher environment (torchrun: RANK / WORLD_SIZE / LOCAL_RANK / MASTER_PORT; mpirun: OMPI_COMM_WORLD_*): one process
    // per GPU of ONE node. Rank 0 publishes the NCCL unique id through a file named after the launcher (parent pid + port),
    // the other ranks poll for it. Selects device LOCAL_RANK. world == 1 (or no launcher) -> *out = NULL, WOST_OK.
    wost_status wost_comm_create_from_env(wost_comm **out)
    {
        if (!out)
            return WOST_ERR_INVALID;
        *out = nullptr;
        auto env_int = [](const char *a, const char *b, int dflt)
        {
            const char *v = getenv(a);
            if (!v || !*v)
                v = getenv(b);
            return (v && *v) ? atoi(v) : dflt;
        };
        const int world = env_int("WORLD_SIZE", "OMPI_COMM_WORLD_SIZE", 1);
        const int rank = env_int("RANK", "OMPI_COMM_WORLD_RANK", 0);
        const int local = env_int("LOCAL_RANK", "OMPI_COMM_WORLD_LOCAL_RANK", rank);
        if (world <= 1)
            return WOST_OK;
        if (cudaSetDevice(local) != cudaSuccess)
        {
            set_last_error("wost_comm_create_from_env: cudaSetDevice(LOCAL_RANK) failed");
            return WOST_ERR_CUDA;
        }
        std::string path;
        if (const char *f = getenv("WOST_COMM_FILE"))
            path = f;
        else
        {
            const char *port = getenv("MASTER_PORT");
            const char *tmp = getenv("TMPDIR");
            path = std::string(tmp && *tmp ? tmp : "/tmp") + "/wost_nccl_id_" + std::to_string((long)getppid()) + "_" + (port ? port : "0");
            // per-launch nonce: a torchrun agent that restarts its workers keeps its pid and port, so a file left behind by a
            // rank 0 that died before the unlink below must not be mistaken for this launch's id
            for (const char *key : {"TORCHELASTIC_RUN_ID", "TORCHELASTIC_RESTART_COUNT", "OMPI_MCA_ess_base_jobid"})
                if (const char *v = getenv(key))
                {
                    path += "_";
                    for (const char *ch = v; *ch && ch - v < 64; ++ch)
                        path += (isalnum((unsigned char)*ch) ? *ch : '-');
                }
        }
        unsigned char id[128];
        if (rank == 0)
        {
            const wost_status rc = wost_comm_unique_id(id);
            if (rc != WOST_OK)
                return rc;
            const std::string tmpf = path + ".tmp";
            // the name is predictable: never write through a link someone else planted, never reuse a file we did not create
            unlink(path.c_str());
            unlink(tmpf.c_str());
            const int fd = open(tmpf.c_str(), O_WRONLY | O_CREAT | O_EXCL | O_NOFOLLOW, 0600);
            if (fd < 0 || write(fd, id, sizeof(id)) != (ssize_t)sizeof(id))
            {
                if (fd >= 0)
                    close(fd);
                set_last_error("wost_comm_create_from_env: cannot write " + tmpf);
                return WOST_ERR_IO;
            }
            close(fd);
            if (rename(tmpf.c_str(), path.c_str()) != 0)
            {
                set_last_error("wost_comm_create_from_env: cannot publish " + path);
                return WOST_ERR_IO;
            }
        }
        else
        {
            bool got = false;
            for (int attempt = 0; attempt < 6000 && !got; ++attempt) // 60 s
            {
                struct stat sb;
                if (lstat(path.c_str(), &sb) == 0 && S_ISREG(sb.st_mode) && sb.st_size == (off_t)sizeof(id))
                {
                    const int fd = open(path.c_str(), O_RDONLY | O_NOFOLLOW);
                    if (fd >= 0)
                    {
                        got = read(fd, id, sizeof(id)) == (ssize_t)sizeof(id);
                        close(fd);
                    }
                }
                if (!got)
                    std::this_thread::sleep_for(std::chrono::milliseconds(10));
            }
            if (!got)
            {
                set_last_error("wost_comm_create_from_env: timed out waiting for " + path);
                return WOST_ERR_IO;
            }
        }
        const wost_status rc = wost_comm_create(rank, world, id, out);
        if (rc == WOST_OK && rank == 0)
            (*out)->id_file = path; // ncclCommInitRank is collective: every rank has read the id by now
        if (rank == 0)
            unlink(path.c_str());
        return rc;
    }

    void wost_comm_destroy(wost_comm *c)
    {
        if (!c)
            return;
        if (c->comm)
            if (NcclApi *api = nccl_api())
                api->CommDestroy(c->comm);
        delete c;
    }

    wost_status wost_comm_info(const wost_comm *c, int *rank, int *world, int *nccl_version)
    {
        if (rank)
            *rank = c ? c->rank : 0;
        if (world)
            *world = c ? c->world : 1;
        if (nccl_version)
        {
            *nccl_version = 0;
            if (NcclApi *api = nccl_api())
                if (api->GetVersion)
                    api->GetVersion(nccl_version);
        }
        return WOST_OK;
    }

    // shard settings of this rank (interleaved row tiles); no-op for a NULL / single-rank communicator
    wost_status wost_comm_shard_settings(const wost_comm *c, wost_settings *s, int stride_rows)
    {
        if (!s)
            return WOST_ERR_INVALID;
        if (!c || c->world <= 1)
            return WOST_OK;
        s->pixel_stride_rows = stride_rows > 0 ? stride_rows : 4;
        s->num_shards = c->world;
        s->shard = c->rank;
        return WOST_OK;
    }

    // guided training: ncclAllReduce(sum) of the gradient slab on the integrator's stream, once per optimiser step
    wost_status wost_integrator_set_comm(wost_integrator *it, wost_comm *c)
    {
        if (!it)
            return WOST_ERR_INVALID;
        if (!c || c->world <= 1)
            return wost_set_allreduce(it, nullptr, nullptr);
        if (it->s.num_shards != c->world || it->s.shard != c->rank)
        {
            set_last_error("wost_integrator_set_comm: the integrator's shard settings do not match the communicator (wost_comm_shard_settings)");
            return WOST_ERR_INVALID;
        }
        return wost_set_allreduce(it, comm_allreduce_cb, c);
    }

    // Final grid gather: every rank contributes the row tiles it evaluated. root >= 0: afterwards rank `root` holds the
    // complete film of `channel` in its device film buffer (read it with wost_film_download); root < 0: every rank does.
    // Moves (N-1)/N of ONE film over NVLink (packed rows, grouped send/recv or one all-gather) -- not an all-reduce of N films.
    wost_status wost_film_gather(wost_integrator *it, wost_comm *c, int channel, int root)
    {
        if (!it || channel < 0 || channel >= WOST_CH_COUNT)
            return WOST_ERR_INVALID;
        if (!c || c->world <= 1)
            return WOST_OK;
        NcclApi *api = nccl_api();
        const int N = c->world, sr = it->s.pixel_stride_rows, H = it->s.frame_h, rowf = it->s.frame_w * 5;
        if (it->s.num_shards != N || it->s.shard != c->rank || sr <= 0 || root >= N)
        {
            set_last_error("wost_film_gather: the integrator is not sharded by interleaved row tiles over this communicator");
            return WOST_ERR_INVALID;
        }
        cudaStream_t st = it->stream;
        float *film = it->film[channel].ptr;
        int max_rows = 0;
        for (int q = 0; q < N; ++q)
            max_rows = std::max(max_rows, owned_row_count(H, sr, N, q));
        const int my_rows = owned_row_count(H, sr, N, c->rank);
        const size_t slot = (size_t)max_rows * rowf;
        if (c->pack.count < slot && c->pack.alloc(slot) != cudaSuccess)
        {
            set_last_error("wost_film_gather: allocation failed");
            return WOST_ERR_CUDA;
        }
        auto rows_kernel = [&](float *packed, int nrows, int shard, int to_packed)
        {
            const long long n = (long long)nrows * rowf;
            if (n > 0)
                k_film_rows<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(film, packed, rowf, nrows, sr, N, shard, to_packed);
        };
        if (root < 0)
        {
            if (c->stage.count < slot * N && c->stage.alloc(slot * N) != cudaSuccess)
            {
                set_last_error("wost_film_gather: allocation failed");
                return WOST_ERR_CUDA;
            }
            rows_kernel(c->pack.ptr, my_rows, c->rank, 1);
            WOST_NCCL_TRY(api->AllGather(c->pack.ptr, c->stage.ptr, slot, ncclFloat, c->comm, st));
            for (int q = 0; q < N; ++q)
                if (q != c->rank)
                    rows_kernel(c->stage.ptr + slot * q, owned_row_count(H, sr, N, q), q, 0);
        }
        else if (c->rank != root)
        {
            rows_kernel(c->pack.ptr, my_rows, c->rank, 1);
            WOST_NCCL_TRY(api->Send(c->pack.ptr, (size_t)my_rows * rowf, ncclFloat, root, c->comm, st));
        }
        else
        {
            if (c->stage.count < slot * N && c->stage.alloc(slot * N) != cudaSuccess)
            {
                set_last_error("wost_film_gather: allocation failed");
                return WOST_ERR_CUDA;
            }
            WOST_NCCL_TRY(api->GroupStart());
            ncclResult_t recv_rc = ncclSuccess; // a failed Recv must not leave the group open: close it, then report
            for (int q = 0; q < N && recv_rc == ncclSuccess; ++q)
                if (q != root)
                    recv_rc = api->Recv(c->stage.ptr + slot * q, (size_t)owned_row_count(H, sr, N, q) * rowf, ncclFloat, q, c->comm, st);
            const ncclResult_t end_rc = api->GroupEnd();
            WOST_NCCL_TRY(recv_rc);
            WOST_NCCL_TRY(end_rc);
            for (int q = 0; q < N; ++q)
                if (q != root)
                    rows_kernel(c->stage.ptr + slot * q, owned_row_count(H, sr, N, q), q, 0);
        }
        WOST_CUDA_TRY(cudaGetLastError());
        WOST_CUDA_TRY(cudaStreamSynchronize(st));
        return WOST_OK;
    }

    // host scalars across ranks (timing: max; counters: sum) -- what elaina-exec needs to report a whole-job result.json
    wost_status wost_comm_allreduce_f64(wost_comm *c, double *vals, int n, int op_max)
    {
        if (!vals || n < 0)
            return WOST_ERR_INVALID;
        if (!c || c->world <= 1 || n == 0)
            return WOST_OK;
        NcclApi *api = nccl_api();
        if (c->scal.count < (size_t)n && c->scal.alloc((size_t)n) != cudaSuccess)
            return WOST_ERR_CUDA;
        WOST_CUDA_TRY(cudaMemcpy(c->scal.ptr, vals, sizeof(double) * n, cudaMemcpyHostToDevice));
        WOST_NCCL_TRY(api->AllReduce(c->scal.ptr, c->scal.ptr, (size_t)n, ncclDouble, op_max ? ncclMax : ncclSum, c->comm, (cudaStream_t)0));
        WOST_CUDA_TRY(cudaStreamSynchronize(0));
        WOST_CUDA_TRY(cudaMemcpy(vals, c->scal.ptr, sizeof(double) * n, cudaMemcpyDeviceToHost));
        return WOST_OK;
    }
}
