/*
 * litap.h — C ABI of liblitap_cuda.so: the B200 (sm_100a) implementation of the raster
 * core of LITAP's flow_mapper() / form_mapper().
 *
 * The reference (FRDC-SHL/LITAP, pure R) has no FFI for this path; the boundary is the set of
 * internal R functions below.  Each entry point replaces one of them and is what the R-side
 * `.Call` shim (litap_b200/r/src/litap_call.c) binds:
 *
 *   litap_ddir   <- calc_ddir2()   R/flow_01_calc_ddir.R:2-118   (called R/flow_mapper.R:157,354)
 *   litap_shed   <- calc_shed4()   R/flow_02_calc_shed.R:10-58   (called R/flow_mapper.R:180,377)
 *                   incl. calc_upslopes(type="upslope")  R/flow_02_calc_shed.R:60-116
 *   litap_form   <- calc_form()    R/form_01_calc_form.R:4-67    (called R/form_mapper.R:103)
 *   litap_weti   <- calc_weti2()   R/form_02b_calc_weti.R:3-110  (called R/form_mapper.R:127)
 *   invert flag  <- invert()       R/flow_06_invert.R:1-5        (called R/flow_mapper.R:343-346)
 *   litap_slope_gc <- slope_gc()   R/flow_05_slope_gc.R:73-164   (called R/flow_mapper.R:307)
 *
 * Grid convention (R/LITAP_load.R:273-332 add_buffer; R/LITAP_functions.R:207-235 nb_values):
 * every array is the BUFFERED grid, NR = nrow+2 by NC = ncol+2, dense row-major, row 0 = north,
 * element i <-> R's seqno = i+1.  The outer ring is NA.  NA elevation = any NaN.
 * Neighbour / direction codes:   7 8 9
 *                                4 5 6      (5 = pit / self, 0 = NA in the uint8 encoding)
 *                                1 2 3
 *
 * Ownership: the caller allocates every host buffer; the library never keeps a caller pointer
 * past return.  Device memory is library-owned and pooled between calls (litap_shutdown frees).
 * Errors: 0 = OK, otherwise a LITAP_E* class; text via litap_last_error() (thread-local).
 * There is NO CPU fallback: without a usable CUDA device every compute call returns LITAP_ECUDA.
 * Threading: calls are synchronous; one host thread at a time per process.
 * Limits: NR*pitch < 2^31 cells per device (32768 x 32768 fits); NC >= 3, NR >= 3.
 */
#ifndef LITAP_H
#define LITAP_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define LITAP_OK 0
#define LITAP_EARG 1   /* bad argument */
#define LITAP_ECUDA 2  /* CUDA runtime / no device */
#define LITAP_ENOMEM 3 /* device or host allocation failed */
#define LITAP_ECOMM 4  /* multi-GPU exchange failed */

/* per-stage device times of the most recent call, milliseconds (CUDA events) */
typedef struct litap_timings {
  float h2d_ms, d2h_ms;
  float ddir_ms;    /* D8 stencil + plateau rounds + flat-pit patches */
  float shed_ms;    /* labels + canonical numbering + upslope + ridge */
  float form_ms;
  float weti_ms;
  float total_ms;   /* whole call, incl. copies */
  int plateau_rounds;
  int weti_queued;  /* calc_weti2: cells handed over through the global ready queues (the others were kept by the lane that released them) */
  int64_t flat_cells, patch_cells, npits;
  int64_t kernel_launches; /* kernels of this library launched by the call */
  float gc_ms;      /* slope_gc */
  float relz_ms;    /* calc_relz + calc_length (device part of litap_relz) */
  float uced_ms;    /* calc_upslopes(type = "uced") */
  int uced_ranges;  /* cell ranges the contributions were processed in (1 unless they exceed the memory budget) */
  float pitstat_ms; /* pit_stat1 (device part of litap_pit_stat) */
  int host_syncs;   /* host waits on the device inside the stages since the last *_flow_form / stage call began */
  double uced_pairs; /* (head, cell below it) contributions = long double sums computed */
} litap_timings;

int litap_init(int device);        /* selects the device and creates the stream; idempotent */
void litap_shutdown(void);         /* frees pooled device memory */
void litap_pool_trim(void);        /* frees the pooled device memory that is not in use by a grid */
const char* litap_last_error(void);
const char* litap_version(void);
int litap_get_timings(litap_timings* out);
/* run on a caller-owned CUDA stream (a cudaStream_t; NULL = back to the library's own stream) */
int litap_set_stream(void* cuda_stream);
/* per-kernel CUDA-event timing, off by default: enable, run stages, then fetch a CSV
 * "kernel,launches,total_ms" (the log is cleared by the report) */
int litap_profile_enable(int on);
int litap_profile_report(char* buf, int64_t cap);

/* ---- host-buffer entry points (what the R shim calls) -------------------------------------- */

/* calc_ddir2.  invert != 0 applies invert() first (elev <- max(elev) - elev) on the device.
 * ddir: 0 = NA, 1..9.  flatcell (may be NULL): 1 = flat after the initial stencil, 0 = not,
 * 2 = NA.  R derives drec = seqno + offset(ddir) (R/LITAP_functions.R:237-254). */
int litap_ddir(const double* elev, int64_t NR, int64_t NC, int invert,
               uint8_t* ddir, uint8_t* flatcell);

/* calc_shed4 (+ calc_upslopes "upslope").  shed: 0 = NA, 1..npits (= initial_shed).
 * upslope: cells draining through the cell incl. itself; 0 where shed is NA.
 * ridge: 1/0, 2 = NA (ring).  Any of shed/upslope/ridge/npits may be NULL. */
int litap_shed(const double* elev, const uint8_t* ddir, int64_t NR, int64_t NC,
               int32_t* shed, uint32_t* upslope, uint8_t* ridge, int64_t* npits);

/* calc_form.  Outputs NaN = NA.  Any output may be NULL. */
int litap_form(const double* elev, int64_t NR, int64_t NC, double grid,
               double* slope_pct, double* slope_deg, double* aspect, double* prof, double* plan);

/* calc_weti2.  upslope as produced by litap_shed (R passes the dem_fill column). */
int litap_weti(const double* elev, const uint8_t* ddir, const uint32_t* upslope,
               int64_t NR, int64_t NC, double grid,
               double* qarea1, double* qarea2, double* qweti1, double* qweti2);

/* calc_upslopes(db, type = "uced")  R/flow_02_calc_shed.R:60-145 (uced_values :118-130), called by pit removal at
 * R/flow_04_remove_pit_fast.R:73,345: upslope cumulative elevation drop.  shed = the shedno column of db (0 = NA:
 * those cells keep NA); ddir as in litap_shed (drec is derived from it, R/LITAP_functions.R:237-254).
 * uced: NaN = NA (cells without a shed, pits nothing flows into).  Bit-identical with the reference's
 * long double sum() per track and its order of additions. */
int litap_uced(const double* elev, const uint8_t* ddir, const int32_t* shed, int64_t NR, int64_t NC, double* uced);

/* pit_stat1(db, shed = ..., method)  R/flow_03_pit_stat_fast.R:1-162 (flow_mapper: R/flow_mapper.R:181,222,378,416; pit
 * removal: R/flow_04_remove_pit_fast.R:19,44,90,136,240,270): the pour point of every shed (lowest ridge cell / neighbour
 * pair that tips into another shed) and the statistics of its pit.  shed = the column named by `shed` (0 = NA); ridge as
 * returned by litap_shed (1 = ridge cell); edge_map 0 / 1; method 0 = "litap", 1 = "landmapr".
 * out[k] (k < 21, may be NULL) receives column litap_pit_stat_names()[k] -- shedno shed_area edge_pit pit_area pit_vol
 * pit_elev pit_seqno pit_row pit_col pre_vol varatio in_seqno in_row in_col in_elev out_seqno out_row out_col out_elev
 * drains_to pour_elev, seqno / row / col in the buffered grid's 1-based numbering, NaN = NA --, one row per shed with
 * a pour point in shedno order, at most cap rows; *nrows = rows of the table.  pit_vol is R's long double sum() in
 * the reference's row order. */
int litap_pit_stat(const double* elev, const uint8_t* ddir, const int32_t* shed, const uint32_t* upslope,
                   const uint8_t* ridge, const uint8_t* edge_map, int64_t NR, int64_t NC, int method,
                   int64_t cap, double* const* out, int64_t* nrows);
const char* const* litap_pit_stat_names(int* n);

/* slope_gc: gradients towards east / north with their zero signs settled, magnitudes (0 -> 0.00001),
 * row / column curvatures, hillslope run labels per row and per column.  f64 outputs NaN = NA;
 * hill_*_dir (2 = east, 4 = west; 1 = north, 3 = south) and hill_*_n / hill_*_cell: 0 = NA.
 * The grid must carry the NA ring of add_buffer().  Any output may be NULL. */
int litap_slope_gc(const double* elev, int64_t NR, int64_t NC, double grid,
                   double* sgre, double* sgr, double* sgcn, double* sgc, double* scr, double* scc,
                   uint8_t* hill_r_dir, uint8_t* hill_c_dir, int32_t* hill_r_n, int32_t* hill_r_cell,
                   int32_t* hill_c_n, int32_t* hill_c_cell);

/* calc_relz + calc_length (R/form_03_calc_relz.R:3-48,194-276,424-514; R/form_04_calc_length.R:3-35; called
 * R/form_mapper.R:154,177) on the buffered grids form_mapper works on: db = the filled DEM of flow_mapper
 * (elev, ddir, drec, upslope, fill_shed, local_shed), idb = the inverted DEM (elev, ddir, drec, upslope,
 * inv_local_shed), the pond table = pit of every local shed (npond = 0: calc_pit3's "ddir == 5 cells" branch,
 * ddir required; otherwise ddir may be NULL).  Integer inputs: 0 = NA; drec is R's 1-based seqno.  Cells
 * without elevation are expected to be NA in every column and not to be drained into, as in LITAP's grids;
 * n2pit / n2peak are reported for cells that have a shed number.
 * out[k], k < LITAP_RELZ_NOUT (any may be NULL): the columns calc_length returns, in the order of
 * litap_relz_names(); NaN = NA; *_row / *_col are buffered 1-based numbers as in R before remove_buffer().
 * Errors: LITAP_EARG for drec / shed numbers outside the grid or flow paths that never end. */
#define LITAP_RELZ_NOUT 39
int litap_relz(int64_t NR, int64_t NC, const double* elev, const uint8_t* ddir, const int32_t* drec,
               const int32_t* upslope, const int32_t* fill_shed, const int32_t* local_shed,
               const double* ielev, const uint8_t* iddir, const int32_t* idrec, const int32_t* iupslope,
               const int32_t* inv_local_shed, int64_t npond, const int32_t* pond_shedno,
               const int32_t* pond_row, const int32_t* pond_col, const double* pond_elev, double str_val,
               double ridge_val, double grid, double* const* out);
const char* litap_relz_names(void);   /* the LITAP_RELZ_NOUT column names, comma-separated */

/* Fused composite (ddir -> shed/upslope/ridge -> form) with one upload of elev; the
 * intermediate ddir never leaves the device.  Any output may be NULL (then not downloaded). */
int litap_flow_form(const double* elev, int64_t NR, int64_t NC, double grid, int invert,
                    uint8_t* ddir, uint8_t* flatcell, int32_t* shed, uint32_t* upslope,
                    uint8_t* ridge, int64_t* npits, double* slope_pct, double* slope_deg,
                    double* aspect, double* prof, double* plan);

/* ---- rasters on disk: ESRI floating-point raster = <name>.flt (float32, row-major, north row first)
 * + <name>.hdr (ncols, nrows, xllcorner, yllcorner, cellsize, NODATA_value, byteorder).  Replaces, for big DEMs, the
 * raster::raster() %>% as.data.frame() path of load_file (R/LITAP_load.R:77-79,187-192) and the .rds / .csv tibbles of
 * save_output (R/LITAP_read_write.R:8-66). */
typedef struct litap_raster_info {
  int64_t nrow, ncol;
  double xll, yll, cellsize, nodata;
  int big_endian;
} litap_raster_info;
typedef struct litap_grid litap_grid;
int litap_flt_info(const char* path /* .flt or .hdr */, litap_raster_info* out);
/* a new grid (NA ring added like add_buffer(), NODATA_value and missing_value -> NA like db_format()) from the file, read
 * in row chunks through pinned buffers: the disk, PCIe and the float32 -> fp64 kernel overlap.  info may be NULL. */
int litap_grid_read_flt(const char* path, double missing_value, litap_grid** out, litap_raster_info* info);
/* one result of the grid without its ring (remove_buffer()) as a headerless little-endian binary file in its own dtype
 * (as_flt = 0: f64 / i32 / u32 / u8; NA = NaN or 0) or narrowed to float32 (as_flt = 1), plus <path minus extension>.hdr.
 * georef (may be NULL) supplies xllcorner / yllcorner / cellsize for the header. */
int litap_grid_write_raw(litap_grid* g, const char* name, const char* path, int as_flt, const litap_raster_info* georef);
/* n results at once, one writer per file (up to 8 side by side): a fresh file fills its page-cache pages at ~4 GB/s however
 * many threads store into it, several files do so independently */
int litap_grid_write_raw_many(litap_grid* g, int n, const char* const* names, const char* const* paths, int as_flt,
                              const litap_raster_info* georef);

/* ---- device-resident entry points (bench / large DEMs that never exist on the host) -------- */

/* litap_grid is opaque: one DEM (or one row band) resident in HBM */

int litap_grid_create(int64_t NR, int64_t NC, litap_grid** out);
void litap_grid_destroy(litap_grid* g);
/* fill elev with the synthetic hash-fBm DEM; rows row0.. of a gnr-row DEM */
int litap_grid_synth(litap_grid* g, int64_t row0, int64_t gnr, uint32_t seed, double quant, int na_permille);
int litap_grid_upload_elev(litap_grid* g, const double* elev);
/* run stages on the resident grid; results stay on the device */
int litap_grid_ddir(litap_grid* g, int invert);
int litap_grid_shed(litap_grid* g);
int litap_grid_form(litap_grid* g, double grid_m);
int litap_grid_weti(litap_grid* g, double grid_m);
int litap_grid_uced(litap_grid* g);
/* pit_stat1 on the shed / upslope / ridge columns litap_grid_shed left on the grid (shed = "initial_shed"); edge_map is
 * derived from the elevations on the device.  Table as for litap_pit_stat. */
int litap_grid_pit_stat(litap_grid* g, int method, int64_t cap, double* const* out, int64_t* nrows);   /* needs ddir and shed on the grid (litap_grid_ddir + litap_grid_shed); result "uced" */
int litap_grid_slope_gc(litap_grid* g, double grid_m);
/* calc_form + calc_ddir2 + calc_shed4 as one call: form overlaps the flow stages on a second stream */
int litap_grid_flow_form(litap_grid* g, double grid_m, int invert);
/* copy results out (dense NR*NC); name is one of: elev ddir flatcell shed upslope ridge
 * slope_pct slope_deg aspect prof plan qarea1 qarea2 qweti1 qweti2
 * sgre sgr sgcn sgc scr scc hill_r_dir hill_c_dir hill_r_n hill_r_cell hill_c_n hill_c_cell */
int litap_grid_download(litap_grid* g, const char* name, void* dst);
int64_t litap_grid_npits(const litap_grid* g);
/* order-independent checksum (sum of value*index hashes) of a result array, computed on the device */
int litap_grid_checksum(litap_grid* g, const char* name, uint64_t* out);
/* conservation check of calc_shed4: number of cells with a shed and the sum of upslope over the pits among them (a band:
 * its own rows); over a whole DEM the two are equal.  litap_bands_shed uses it to refuse DEMs in which a 32-bit count wrapped. */
int litap_grid_conservation(litap_grid* g, uint64_t* cells, uint64_t* pit_upslope);

/* litap_checksum_noalloc has no counterpart in the reference: it was added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, serving loops, harnesses that forbid it).
 * The checksum of litap_grid_checksum over a device array the CALLER owns: NR rows x NC columns of elem_size (1, 4 or 8)
 * bytes, `pitch` elements from one row to the next, result into the device word dev_out, both kernels enqueued on
 * cuda_stream (a cudaStream_t; NULL = the default stream) with no synchronisation.  Calls no cudaMalloc / cudaFree and
 * no allocating library routine: its scratch (one partial sum per raster row) is a static __device__ array of
 * LITAP_NOALLOC_MAX_ROWS words (512 KB) that every GPU which loads the library holds until exit.  NR beyond that
 * limit is refused with LITAP_EARG.  All calls share the array: calls that overlap on the GPU race (order them on one
 * stream).  The first call on a device looks the array's address up (cudaGetSymbolAddress); make it outside capture.
 * Works on the current device; needs no litap_init. */
#define LITAP_NOALLOC_MAX_ROWS 65536
int litap_checksum_noalloc(const void* dev_arr, int elem_size, int64_t NR, int64_t NC, int64_t pitch,
                           uint64_t* dev_out, void* cuda_stream);

/* ---- row bands: one DEM sharded over several GPUs ----------------------- */
/* The DEM (gnr rows x nc cols, unbuffered) is cut into nbands contiguous row bands; band b owns rows
 * [gnr*b/nbands, gnr*(b+1)/nbands) and keeps one halo row on either side.  A band is a litap_grid.
 * Two ways to run:
 *   - one band per process / GPU (torchrun): rank 0 calls litap_comm_unique_id, the 128 bytes are
 *     broadcast by the host (torch.distributed, MPI, a file ...), every rank calls litap_comm_init
 *     and passes its single band to the litap_bands_* calls (collective: all ranks must call);
 *   - all bands in one process on one device, no communicator: pass all nbands bands in band order
 *     (used to verify that the sharded result equals the unsharded one bit for bit).
 * calc_weti2 is not sharded (its visiting order couples all bands): run it on a whole-DEM grid. */
int litap_comm_unique_id(void* id128);
int litap_comm_init(int rank, int nranks, const void* id128);
void litap_comm_destroy(void);
int litap_band_create(int64_t gnr, int64_t nc, int band, int nbands, litap_grid** out);
int litap_band_synth(litap_grid* band, uint32_t seed, double quant, int na_permille);
/* elev_full: the whole buffered DEM, (gnr+2) x (nc+2) dense; the band copies its rows and halos */
int litap_band_upload_elev(litap_grid* band, const double* elev_full);
/* rows: only this band's rows of the buffered DEM, halo rows included: (own rows + 2) x (nc+2) dense */
int litap_band_upload_rows(litap_grid* band, const double* rows);
int litap_bands_ddir(litap_grid** bands, int n, int invert);
int litap_bands_shed(litap_grid** bands, int n);
int litap_bands_form(litap_grid** bands, int n, double grid_m);
/* form + ddir + shed of the local bands as one call (calc_form overlaps the flow stages, see litap_grid_flow_form) */
int litap_bands_flow_form(litap_grid** bands, int n, double grid_m, int invert);
/* the rows of the full buffered DEM this band answers for (own rows; + the NA ring row for the
 * first / last band): first row and row count; litap_band_download copies exactly those rows */
int litap_band_rows(const litap_grid* band, int64_t* row0, int64_t* nrows);
int litap_band_download(litap_grid* band, const char* name, void* dst);
/* the band's share of litap_grid_checksum of the whole DEM (same rows as litap_band_download, hashed with their cell
 * numbers in the whole DEM): the wrapping sum over all bands equals the checksum of the unsharded result */
int litap_band_checksum(litap_grid* band, const char* name, uint64_t* out);

#ifdef __cplusplus
}
#endif
#endif
