// acino_sba_schur_api.inl — drivers of the points+extrinsics solver (included by acino.cu after the handle).
namespace acino {
static int schur_reset(acino_sba*) { return ACINO_OK; }
static int solve_schur(acino_sba*, const acino_sba_opts*, acino_sba_stats*) { set_error("points+extrinsics mode not built yet"); return ACINO_ESTATE; }
static int schur_iterate(acino_sba*, const acino_sba_opts*) { set_error("points+extrinsics mode not built yet"); return ACINO_ESTATE; }
static void comm_destroy(acino_sba*) {}
}
extern "C" {
int acino_sba_get_rvec_t(acino_sba_t*, double*, double*) { set_error("not built yet"); return ACINO_ESTATE; }
int acino_sba_set_rvec_t(acino_sba_t*, const double*, const double*) { set_error("not built yet"); return ACINO_ESTATE; }
int acino_nccl_unique_id(void*) { set_error("not built yet"); return ACINO_ENCCL; }
int acino_sba_comm_init(acino_sba_t*, const void*, int, int) { set_error("not built yet"); return ACINO_ENCCL; }
}
