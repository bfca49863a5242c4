// nccl_dyn.cpp -- run-time binding of NCCL (see nccl_dyn.h).
#include "nccl_dyn.h"

#include <dlfcn.h>
#include <link.h>

#include <cstdlib>
#include <mutex>

namespace paths {

namespace {
NcclApi g_api;
bool g_tried = false, g_ok = false;
std::string g_why;
std::mutex g_mu;

template <typename F> bool bind(void *h, const char *name, F &out, std::string &err) {
    void *p = dlsym(h, name);
    if (!p) {
        err = std::string("NCCL library lacks ") + name;
        return false;
    }
    out = reinterpret_cast<F>(p);
    return true;
}
} // namespace

const NcclApi *nccl_api(std::string &err) {
    std::lock_guard<std::mutex> lock(g_mu);
    if (!g_tried) {
        g_tried = true;
        void *h = nullptr;
        const char *names[] = {"libnccl.so.2", "libnccl.so"};
        if (const char *forced = getenv("PATHS_NCCL_LIB")) h = dlopen(forced, RTLD_NOW | RTLD_GLOBAL); // a deployment's own choice of file
        for (const char *n : names) // the copy the process already uses, if any
            if (!h) h = dlopen(n, RTLD_NOW | RTLD_NOLOAD);
        for (const char *n : names)
            if (!h) h = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
        if (!h) {
            g_why = std::string("cannot load libnccl.so.2: ") + (dlerror() ? dlerror() : "not found");
        } else {
            bool ok = bind(h, "ncclGetVersion", g_api.GetVersion, g_why) && bind(h, "ncclGetUniqueId", g_api.GetUniqueId, g_why) &&
                      bind(h, "ncclCommInitRank", g_api.CommInitRank, g_why) && bind(h, "ncclCommInitAll", g_api.CommInitAll, g_why) &&
                      bind(h, "ncclCommDestroy", g_api.CommDestroy, g_why) && bind(h, "ncclReduce", g_api.Reduce, g_why) &&
                      bind(h, "ncclSend", g_api.Send, g_why) && bind(h, "ncclRecv", g_api.Recv, g_why) &&
                      bind(h, "ncclGroupStart", g_api.GroupStart, g_why) && bind(h, "ncclGroupEnd", g_api.GroupEnd, g_why) &&
                      bind(h, "ncclGetErrorString", g_api.GetErrorString, g_why);
            if (ok) {
                struct link_map *lm = nullptr;
                if (dlinfo(h, RTLD_DI_LINKMAP, &lm) == 0 && lm && lm->l_name) g_api.origin = lm->l_name;
                g_ok = true;
            }
        }
    }
    if (!g_ok) {
        err = g_why;
        return nullptr;
    }
    return &g_api;
}

} // namespace paths
