#include "rgbdfe_host.h"

#include <cstdio>
#include <cstdlib>
#include <mutex>
#include <vector>

namespace rgbdfe_host {

namespace {
    int g_w = 1920, g_h = 1080, g_feats = 4000, g_device = 0;
    std::mutex g_mutex;
    bool g_seeded = false;
    unsigned g_seed = 0;

    struct Holder {
        rgbdfe_ctx* c = nullptr;
        ~Holder()
        {
            if (c)
                rgbdfe_destroy(c);
        }
    };
    thread_local Holder t_holder;
}

void set_capacity(int max_w, int max_h, int max_feats)
{
    std::lock_guard<std::mutex> lk(g_mutex);
    g_w = max_w, g_h = max_h, g_feats = max_feats;
}

void set_device(int device)
{
    std::lock_guard<std::mutex> lk(g_mutex);
    g_device = device;
}

rgbdfe_ctx* ctx()
{
    if (!t_holder.c) {
        std::lock_guard<std::mutex> lk(g_mutex);
        rgbdfe_ctx* c = nullptr;
        const int rc = rgbdfe_create(&c, g_device, g_w, g_h, g_feats, 1);
        if (rc != RGBDFE_OK || !c) {
            fprintf(stderr, "rgbdfe: cannot create a CUDA context (code %d): %s -- there is no CPU fallback\n", rc,
                c ? rgbdfe_last_error(c) : "no device");
            std::terminate();
        }
        if (g_seeded)
            rgbdfe_rng_seed(c, g_seed);
        t_holder.c = c;
    }
    return t_holder.c;
}

const void*& config_owner()
{
    thread_local const void* owner = nullptr;
    return owner;
}

int report(const char* where, int rc)
{
    if (rc != RGBDFE_OK)
        fprintf(stderr, "rgbdfe: %s failed (code %d): %s\n", where, rc, t_holder.c ? rgbdfe_last_error(t_holder.c) : "");
    return rc;
}

bool seeded()
{
    std::lock_guard<std::mutex> lk(g_mutex);
    return g_seeded;
}

void seed(unsigned s)
{
    {
        std::lock_guard<std::mutex> lk(g_mutex);
        g_seeded = true;
        g_seed = s;
    }
    report("rgbdfe_rng_seed", rgbdfe_rng_seed(ctx(), s));
}

} // namespace rgbdfe_host
