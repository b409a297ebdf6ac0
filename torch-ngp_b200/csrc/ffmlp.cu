// ffmlp.cu -- placeholder until the tcgen05 fused MLP lands; fails loudly.
#include "common.cuh"
extern "C" {
int ngp_ffmlp_forward(const void *, const void *, int, uint32_t, uint32_t, uint32_t, uint32_t, uint32_t, uint32_t, uint32_t, void *, void *, void *) {
    ngp::set_error("ffmlp_forward: not implemented yet"); return NGP_EUNSUPPORTED; }
int ngp_ffmlp_inference(const void *, const void *, int, uint32_t, uint32_t, uint32_t, uint32_t, uint32_t, uint32_t, uint32_t, void *, void *, void *) {
    ngp::set_error("ffmlp_inference: not implemented yet"); return NGP_EUNSUPPORTED; }
int ngp_ffmlp_backward(const void *, const void *, const void *, int, const void *, uint32_t, uint32_t, uint32_t, uint32_t, uint32_t, uint32_t, uint32_t, int,
                       void *, void *, void *, int, void *, size_t, void *) {
    ngp::set_error("ffmlp_backward: not implemented yet"); return NGP_EUNSUPPORTED; }
size_t ngp_ffmlp_backward_workspace_bytes(uint32_t, uint32_t, uint32_t, uint32_t) { return 0; }
}
