// scenegen_stub.cpp — the two error-reporting entry points of crt_api.cu, for libcrt_scenegen.so: the HOST-ONLY library that
// holds the scene generators (scenes.cpp) and the image decoder (jpeg_decode.cpp) and nothing that touches a GPU. It exists so
// that a process which only needs a benchmark scene description — bench.py --impl reference, the CPU baseline of the
// reference's own code — does not map the product library ("the product library has no business in that process").
#include "crt_internal.h"

#include <mutex>

static std::mutex g_mu;
static std::string g_err;

crt_status crt_set_error(crt_context*, crt_status st, const char* msg)
{
    std::lock_guard<std::mutex> lk(g_mu);
    g_err = msg ? msg : "";
    return st;
}

extern "C" const char* crt_last_error_string(const crt_context*)
{
    std::lock_guard<std::mutex> lk(g_mu);
    static thread_local std::string copy;
    copy = g_err;
    return copy.c_str();
}
