// tp.cpp -- NCCL entry points resolved at run time.  The engine has no link-time NCCL dependency:
// a host process that already loaded NCCL (PyTorch bundles its own copy) keeps using that copy,
// and single-GPU engines never touch NCCL at all.
#include <dlfcn.h>

#include "engine.h"

namespace dis {

NcclApi *nccl_api(std::string *err) {
  static NcclApi api;
  static bool tried = false, ok = false;
  static std::string saved;
  if (!tried) {
    tried = true;
    void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) { saved = std::string("cannot load libnccl.so.2: ") + dlerror(); }
    else {
      api.lib = h;
      api.GetUniqueId = reinterpret_cast<decltype(api.GetUniqueId)>(dlsym(h, "ncclGetUniqueId"));
      api.CommInitRank = reinterpret_cast<decltype(api.CommInitRank)>(dlsym(h, "ncclCommInitRank"));
      api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(dlsym(h, "ncclCommDestroy"));
      api.AllReduce = reinterpret_cast<decltype(api.AllReduce)>(dlsym(h, "ncclAllReduce"));
      api.AllGather = reinterpret_cast<decltype(api.AllGather)>(dlsym(h, "ncclAllGather"));
      api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(dlsym(h, "ncclGetErrorString"));
      ok = api.GetUniqueId && api.CommInitRank && api.CommDestroy && api.AllReduce && api.AllGather && api.GetErrorString;
      if (!ok) saved = "libnccl.so.2 lacks an expected symbol";
    }
  }
  if (!ok) { if (err) *err = saved; return nullptr; }
  return &api;
}

}  // namespace dis
