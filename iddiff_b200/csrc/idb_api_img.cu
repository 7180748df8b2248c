// C ABI, image-network side (U-Net score models): see include/iddiff_b200.h.
#include "../../include/iddiff_b200.h"
#include "idb_kernels.h"

using idb::api_check;
using idb::api_fail;

extern "C" {

int idb_upfirdn2d(const float* input, const float* kernel, int major, int in_h, int in_w, int minor, int kernel_h,
                  int kernel_w, int up_x, int up_y, int down_x, int down_y, int pad_x0, int pad_x1, int pad_y0,
                  int pad_y1, float* out, void* stream) {
    if (!input || !kernel || !out) return api_fail("idb_upfirdn2d: null pointer");
    if (major < 0 || in_h <= 0 || in_w <= 0 || minor <= 0 || kernel_h <= 0 || kernel_w <= 0)
        return api_fail("idb_upfirdn2d: bad shape");
    if (up_x <= 0 || up_y <= 0 || down_x <= 0 || down_y <= 0) return api_fail("idb_upfirdn2d: up/down must be >= 1");
    if (kernel_h * kernel_w > 1024) return api_fail("idb_upfirdn2d: FIR kernel larger than 1024 taps");
    idb::UpFirDnParams p;
    p.in = input; p.kernel = kernel; p.out = out;
    p.major = major; p.in_h = in_h; p.in_w = in_w; p.minor = minor; p.kh = kernel_h; p.kw = kernel_w;
    p.up_x = up_x; p.up_y = up_y; p.down_x = down_x; p.down_y = down_y;
    p.pad_x0 = pad_x0; p.pad_x1 = pad_x1; p.pad_y0 = pad_y0; p.pad_y1 = pad_y1;
    p.out_h = p.out_w = p.minor_tile = 0;
    return api_check(idb::launch_upfirdn2d(p, idb::api_num_sms(), static_cast<cudaStream_t>(stream)), "idb_upfirdn2d");
}

int idb_fused_bias_act(const float* input, const float* bias, const float* refer, int act, int grad, float alpha,
                       float scale, long long size_x, int step_b, int size_b, float* out, void* stream) {
    if (!input || !out || size_x < 0) return api_fail("idb_fused_bias_act: bad argument");
    if (act != 1 && act != 3) return api_fail("idb_fused_bias_act: act must be 1 (linear) or 3 (leaky ReLU)");
    if (grad < 0 || grad > 2) return api_fail("idb_fused_bias_act: grad must be 0, 1 or 2");
    if (bias && (step_b <= 0 || size_b <= 0)) return api_fail("idb_fused_bias_act: bias needs step_b, size_b > 0");
    return api_check(idb::launch_fused_bias_act(input, bias, refer, act, grad, alpha, scale, size_x, step_b, size_b, out,
                                                idb::api_num_sms(), static_cast<cudaStream_t>(stream)),
                     "idb_fused_bias_act");
}

}  // extern "C"
