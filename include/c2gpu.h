/*
 * c2gpu.h -- C ABI of the B200 (sm_100a) grid field-evaluation module for critic2.
 *
 * This is the drop-in boundary: the entry points a Fortran `interface ... bind(c)` block in critic2 binds
 * (fortran/gpu_interface.f90 shows the binding; INTEGRATION.md the call-site patches).  The reference has no FFI for
 * this path, so the boundary is lifted to the grid loops that call field%grd point by point; each entry cites the
 * reference loop or routine it replaces.  All citations are relative to /root/reference/.
 *
 * Conventions
 *   - every function returns int: 0 ok, -1 "unsupported input" (caller falls through to its own CPU loop),
 *     >0 error (message via c2gpu_last_error).  Nothing aborts or throws across the ABI
 *     (reference error convention: ferror, src/tools_io@proc.F90:1233).
 *   - all coordinates are Cartesian bohr in the cell frame; 3x3 matrices are column-major (Fortran order);
 *     atom / primitive ids are 1-based like the reference.
 *   - the caller owns every host array; *_create copies its inputs to every selected device.
 *   - *_dev variants take DEVICE pointers (on the handle's first device) and a cudaStream_t passed as void*;
 *     they are what a one-process-per-GPU host (bench.py under torchrun) uses.  The plain variants take HOST
 *     pointers, spread the slowest grid index over all devices of the handle and copy the result back.
 *   - there is no CPU fallback inside this library.
 */
#ifndef C2GPU_H
#define C2GPU_H

#ifdef __cplusplus
extern "C" {
#endif

/* ---- library / device management ---------------------------------------------------------------------------- */
/* Select devices 0..min(ndev_req,visible)-1 (ndev_req<=0: all visible).  *ndev = number selected. */
int c2gpu_init(int ndev_req, int *ndev);
/* Select an explicit device list (one-process-per-GPU hosts pass their LOCAL_RANK). */
int c2gpu_init_devices(const int *devs, int n);
int c2gpu_finalize(void);
/* Copy the last error message (NUL terminated) into buf. */
int c2gpu_last_error(char *buf, int len);
/* Number of kernels launched by this library since load (the bench's gpu_launches claim). */
long c2gpu_launch_count(void);

/* ---- field handles ------------------------------------------------------------------------------------------ */
/*
 * Wavefunction field = the flat image of type molwfn after read_* + prune_primitives + complete_struct
 * (src/wfn_private.f90:83-131; tables from src/wfn_private@proc.F90:3077-3184).  Replaces the data side of
 * molwfn%rho2 (src/wfn_private@proc.F90:2032-2228) and calculate_mo_gto (:2673-2781) / calculate_edf (:2783-2873).
 *   xyz        3*natom   env%at(i)%r (cell frame)
 *   icenter    npri      1-based atom of each primitive;  itype npri  primitive type 1..56 (:2695-2705)
 *   expo,dran  npri      exponent and screening radius sqrt(-ln(1e-12)/expo) (:3102-3106)
 *   cmo        column-major ldc x npri (ldc = nmoall >= nmoocc), MO index fastest: cmo(imo,ipri)
 *   occ        nmoocc;   wfntyp 0 rhf, 1 uhf, 2 frac (src/wfn_private.f90:142-144); nalpha: # alpha MOs (uhf)
 *   EDF arrays (nedf may be 0) as read_wfx :864-873; dran_edf like dran.
 *   cell: ismolecule and the matrices field%grd uses to wrap the point (src/fieldmod@proc.f90:698-721).
 *         Only ismolecule != 0 is supported for wavefunction fields (returns -1 otherwise).
 */
int c2gpu_wfn_create(int natom, const double *xyz, int npri, const int *icenter, const int *itype,
                     const double *expo, const double *dran, int nmoocc, int ldc, const double *cmo,
                     const double *occ, int wfntyp, int nalpha, int nedf, const int *icenter_edf,
                     const int *itype_edf, const double *e_edf, const double *c_edf, const double *dran_edf,
                     int ismolecule, const double *m_x2c, const double *m_c2x, void **handle);

/*
 * Promolecular field = environ atoms (cell atoms + periodic images, src/environmod@proc.f90:139-213) + the
 * radial tables agrid(:) (src/grid1mod.f90:25-46).  Replaces environ%promolecular (src/environmod@proc.f90:1202-1327)
 * with grid1%interp (src/grid1mod@proc.f90:84-135).
 *   xyz_env 3*nenv Cartesian (main reduced cell + images); ispc nenv 1-based species; cidx nenv 1-based cell atom
 *   per species: zspc, ngrid, a, b, cutrad (src/global.f90:51-93); tables packed one species after another
 *   (offset of species s = sum ngrid[0..s-1]): tab_r, tab_f, tab_fp, tab_fpp.  ngrid[s] = 0: table not initialised.
 *   m_* : cell and reduced-cell matrices (grd wrap :698-721; y2z_center src/environmod@proc.f90:426-444)
 *   fragmask: NULL or ncell flags (type_promol_frag, :1287-1290)
 */
int c2gpu_promol_create(int nenv, int ncell, const double *xyz_env, const int *ispc, const int *cidx, int nspc,
                        const int *zspc, const int *ngrid, const double *a, const double *b, const double *cutrad,
                        const double *tab_r, const double *tab_f, const double *tab_fp, const double *tab_fpp,
                        int ismolecule, const double *m_x2c, const double *m_c2x, const double *m_xr2c,
                        const double *m_c2xr, const unsigned char *fragmask, void **handle);
int c2gpu_destroy(void *handle);
/*
 * Core augmentation of a field loaded with the CORE / ZPSP options (f%usecore, f%zpsp: src/fieldmod.f90:64,79,
 * src/fieldmod@proc.f90:154-157,184-199): grd adds the promolecular CORE density and its derivatives, evaluated at wc
 * from the tables cgrid(iz,zpsp(iz)) (src/grid1mod@proc.f90:139-164), before the nucleus test / gfmod / del2f
 * (src/fieldmod@proc.f90:827-832; grd0 :922-925).  hcore is a promolecular handle made by c2gpu_promol_create over the
 * same environment with the core tables (ngrid[s] = 0 for species with zpsp <= 0 or Z - zpsp <= 0, which contribute
 * nothing: src/environmod@proc.f90:1299, src/grid1mod@proc.f90:100); NULL detaches.  The core handle is not owned by the
 * field: destroy it after the field.  Every driver of this header (CUBE, NCIPLOT, point lists) then includes the core.
 */
int c2gpu_field_set_core(void *handle, void *hcore);

/* ---- CUBE: rhoplot_cube evaluation loop, src/rhoplot@proc.f90:582-631 ---------------------------------------- */
/* property bits of outmask = nti of :488-523: 0 f, 1 gx, 2 gy, 3 gz, 4 gmod, 5 hxx, 6 hxy, 7 hxz, 8 hyy, 9 hyz,
   10 hzz, 11 lap.  out[] holds one nn1*nn2*nn3 array per set bit (ascending bit order), x fastest (lf(ix,iy,iz)).
   xd: column-major step vectors.  nder is 2 in the reference's loop (:595); 0/1 are accepted (grd semantics).
   The host-array entries (this one, c2gpu_nci_grid, c2gpu_promol_grid) keep their device output arrays in the handle
   between calls (the reference issues one CUBE per property on the same lattice); c2gpu_destroy frees them. */
#define C2GPU_NPROP 12
int c2gpu_eval_grid(void *h, const double x0[3], const double xd[9], const int nn[3], int nder, int periodic,
                    unsigned outmask, double *const *out);
/* slab [iz0,iz1) of the same loop into full-size DEVICE arrays */
/* The *_dev entries run on the handle's first device and share that replica's work counter and scratch buffers: issue the
   calls of ONE handle on ONE stream (or serialise them); different handles may run concurrently on different streams. */
int c2gpu_eval_grid_dev(void *h, const double x0[3], const double xd[9], const int nn[3], int nder, int periodic,
                        unsigned outmask, double *const *out_dev, int iz0, int iz1, void *stream);

/* the layers of height 8 along z with (layer index mod nparts == part), into full-size DEVICE arrays: the
   interleaved (load-balanced) slab decomposition the host variant uses across devices, for one-process-per-GPU hosts */
int c2gpu_eval_grid_dev_part(void *h, const double x0[3], const double xd[9], const int nn[3], int nder, int periodic,
                             unsigned outmask, double *const *out_dev, int part, int nparts, void *stream);
/* height of those layers for handle h (8 for wavefunction fields, 4 for promolecular fields) */
/* Same decomposition one level finer, for the one-process-per-GPU model whose gather is a reduce over full arrays: the
   interleaved unit is one row of super-tiles of a layer (layer_height planes x layer_height lines along y), unit index
   = layer * ceil(nn2 / layer_height) + row, dealt round-robin like the layers.  50 layers over 8 GPUs leave 7 against
   6.25 layers per GPU; 2 500 rows leave 313 against 312.5. */
int c2gpu_eval_grid_dev_part_rows(void *h, const double x0[3], const double xd[9], const int nn[3], int nder,
                                  int periodic, unsigned outmask, double *const *out_dev, int part, int nparts,
                                  void *stream);
int c2gpu_layer_height(void *h);
/* Gather leg of that model ("each GPU D2H-copies straight into its slice of the pinned host array"): the
   blocks part `part` owns (rows = 0: layers, the decomposition of c2gpu_eval_grid_dev_part / c2gpu_nci_grid_dev_part;
   rows = 1: rows of super-tiles, c2gpu_eval_grid_dev_part_rows) go from the part's full-size device arrays into the same
   places of full-size host arrays; nci_order = 1 for crho/cgrad/rhoat (z fastest, x-layers).  With one shared page-locked
   host buffer all ranks copy at once over their own PCIe links; pageable destinations are staged through the handle's
   pinned bounce buffers.  Ordered after the work queued on `stream`; synchronous on return.  Replaces the writer-rank
   gather the reference does not need (its arrays are already in the one process: rhoplot@proc.f90:582, nci@proc.f90:458). */
int c2gpu_gather_part_host(void *h, const int nn[3], int nci_order, int narr, double *const *dev_arr,
                           double *const *host_arr, int part, int nparts, int rows, void *stream);
/* The same two steps on a slab [iz0, iz1) of the slowest output axis (units counted from iz0; iz0 a multiple of the layer
   height): a pass cut into 2-3 slabs lets the gather of slab s run behind the kernel of slab s + 1.  async_ = 1 queues the
   copies behind the work already on `stream` and returns at once (page-locked host arrays only); c2gpu_gather_wait blocks
   until every queued gather of the handle has landed. */
int c2gpu_eval_grid_dev_part_range(void *h, const double x0[3], const double xd[9], const int nn[3], int nder,
                                   int periodic, unsigned outmask, double *const *out_dev, int part, int nparts, int rows,
                                   int iz0, int iz1, void *stream);
int c2gpu_gather_part_host_range(void *h, const int nn[3], int nci_order, int narr, double *const *dev_arr,
                                 double *const *host_arr, int part, int nparts, int rows, int c0, int c1, int async_,
                                 void *stream);
int c2gpu_gather_wait(void *h);

/* ---- NCIPLOT: evaluation loop src/nci@proc.f90:499-561, cutoff pass :569-604 ---------------------------------- */
/* crho/cgrad/rhoat are (0:n3-1,0:n2-1,0:n1-1): z fastest, x slowest (:458-469).  hpromol/rhoat may be NULL. */
int c2gpu_nci_grid(void *href, void *hpromol, const double x0[3], const double xmat[9], const int nstep[3],
                   double *crho, double *cgrad, double *rhoat);
/* slab [i0,i1) of the slowest (x) index into full-size DEVICE arrays */
int c2gpu_nci_grid_dev(void *href, void *hpromol, const double x0[3], const double xmat[9], const int nstep[3],
                       double *crho_dev, double *cgrad_dev, double *rhoat_dev, int i0, int i1, void *stream);
/* interleaved layers along x (see c2gpu_eval_grid_dev_part) */
int c2gpu_nci_grid_dev_part(void *href, void *hpromol, const double x0[3], const double xmat[9], const int nstep[3],
                            double *crho_dev, double *cgrad_dev, double *rhoat_dev, int part, int nparts,
                            void *stream);
/* Cutoff pass (nfrag = 0).  Modifies cgrad in place; sums[3] = sumq,sumqp,sumv (scaled by omega/N, :600-603);
   *ndat = number of (rho,s) pairs of the .dat file; dat (2*maxdat doubles, may be NULL) receives the first maxdat
   pairs in the reference's loop order.  rho_void <= 0 disables the VOID test (rhoat may then be NULL). */
int c2gpu_nci_cutoffs(long npts, double *crho, double *cgrad, const double *rhoat, double rhocut, double dimcut,
                      double rhoplot, const double srhorange[2], int onlyneg, double rho_void, double omega,
                      double sums[3], long *ndat, double *dat, long maxdat);
int c2gpu_nci_cutoffs_dev(long npts, const double *crho_dev, double *cgrad_dev, const double *rhoat_dev,
                          double rhocut, double dimcut, double rhoplot, const double srhorange[2], int onlyneg,
                          double rho_void, double raw_sums[3], long *ndat, double *dat_dev, long maxdat,
                          void *stream);
/* Allocation-free variant of c2gpu_nci_cutoffs_dev.  NO COUNTERPART IN THE REFERENCE: added
   for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
   forbid it).  Same arguments, same kernels and bit-identical results; the pass's scratch (per-block counts, partial sums
   and offsets: 36 bytes per 2 048 points) lives in one static __device__ array of 2 359 552 bytes that every GPU loading
   the library holds until exit, so the entry calls no cudaMalloc / cudaFree / allocating routine.  Limit: npts <=
   C2GPU_NOALLOC_MAX_POINTS (512^3); larger grids are rejected with status 1 and a message (use c2gpu_nci_cutoffs_dev).
   Runs on the CURRENT device.  All calls share the array (it keeps the last call's contents; every entry a call reads is
   rewritten by that call first, so nothing needs clearing): calls that overlap on one GPU race -- serialise them. */
#define C2GPU_NOALLOC_MAX_POINTS 134217728L
int c2gpu_nci_cutoffs_dev_noalloc(long npts, const double *crho_dev, double *cgrad_dev, const double *rhoat_dev,
                                  double rhocut, double dimcut, double rhoplot, const double srhorange[2], int onlyneg,
                                  double rho_void, double raw_sums[3], long *ndat, double *dat_dev, long maxdat,
                                  void *stream);

/* ---- NCIPLOT fragments: src/nci@proc.f90:534-550 (rhofrag) and :579-582 (the `inter` test) ----------------------
   A fragment handle is a promolecular handle made by c2gpu_promol_create over the fragment's own atoms only
   (nenv = ncell = fr%nat, cidx = 1..nat, no periodic images), with cutrad = 1e30 for every species (the reference
   cuts at agrid(iz)%rmax only, :542), ismolecule = 1 and identity cell matrices (the point is not folded, :538).
   rhofrag: [nfrag][n1*n2*n3] = the reference's rhofrag(k,j,i,ifr), z fastest. */
int c2gpu_nci_fragments(int nfrag, void *const *hfrag, const double x0[3], const double xmat[9], const int nstep[3],
                        double *rhofrag);
int c2gpu_nci_fragments_dev(int nfrag, void *const *hfrag, const double x0[3], const double xmat[9], const int nstep[3],
                            double *rhofrag_dev, int i0, int i1, void *stream);
/* the interleaved x-layers (height layer_height = c2gpu_layer_height(href)) of c2gpu_nci_grid_dev_part */
int c2gpu_nci_fragments_dev_part(int nfrag, void *const *hfrag, const double x0[3], const double xmat[9],
                                 const int nstep[3], double *rhofrag_dev, int part, int nparts, int layer_height,
                                 void *stream);
/* Cutoff pass with fragments (nfrag > 0 requires rhofrag and rhoat; defaults rhoparam 0.95, rhoparam2 0.75, :120-121) */
int c2gpu_nci_cutoffs_frag(long npts, double *crho, double *cgrad, const double *rhoat, int nfrag, const double *rhofrag,
                           double rhoparam, double rhoparam2, double rhocut, double dimcut, double rhoplot,
                           const double srhorange[2], int onlyneg, double rho_void, double omega, double sums[3],
                           long *ndat, double *dat, long maxdat);
int c2gpu_nci_cutoffs_frag_dev(long npts, const double *crho_dev, double *cgrad_dev, const double *rhoat_dev, int nfrag,
                               const double *rhofrag_dev, double rhoparam, double rhoparam2, double rhocut,
                               double dimcut, double rhoplot, const double srhorange[2], int onlyneg, double rho_void,
                               double raw_sums[3], long *ndat, double *dat_dev, long maxdat, void *stream);

/* ---- text bodies of the output files: the formatted writes that follow the evaluation ----
   writegrid_cube src/crystalmod@proc.f90:5026-5036, write_cube_body src/nci@proc.f90:867-883, .dat src/nci@proc.f90:593-595.
   fmt   0 '(6(1x,e22.14))' (PRECISECUBE, global@proc.f90:90)   1 '(1p,6(1x,e12.5))'   2 '(6(1x,1p,e13.5e3))' (NCIPLOT)
   order 0 g(n1,n2,n3) x fastest (lf of rhoplot_cube): one record per (ix,iy), iz running   -- writegrid_cube
         1 c(0:n3-1,0:n2-1,0:n1-1) z fastest (crho/cgrad): one record per (i,j), k running  -- write_cube_body
   The text is exactly what the Fortran formatted write produces (6 fields per line, a new line at the end of every
   record, round-half-even decimal conversion); the caller writes it after the header with stream access. */
long c2gpu_cube_text_size(const int n[3], int fmt);
int c2gpu_cube_text(const double *g, const int n[3], int fmt, int order, char *text);
int c2gpu_cube_text_dev(const double *g_dev, const int n[3], int fmt, int order, char *text_dev /*16-byte aligned*/,
                        void *stream);
/* '(1p,E15.7,E15.7)' lines of the .dat file from the (rho,s) pairs of c2gpu_nci_cutoffs: 31 bytes per pair */
long c2gpu_dat_text_size(long npairs);
int c2gpu_dat_text(long npairs, const double *dat, char *text);

/* ---- whole-cell promolecular grid: crystal%promolecular_grid, src/crystalmod@proc.f90:5118-5158 --------------- */
int c2gpu_promol_grid(void *hpromol, const int n[3], double *out);
int c2gpu_promol_grid_dev(void *hpromol, const int n[3], double *out_dev, int iz0, int iz1, void *stream);

/* ---- arbitrary point lists: field%grd, src/fieldmod@proc.f90:613-842 ("next" row: mesh%fill, planes, lines) ---- */
/* res: npts x C2GPU_RES_LEN doubles: [0] f, [1..3] gf, [4..12] hf (column-major), [13] gfmod, [14] del2f, [15] isnuc */
#define C2GPU_RES_LEN 16
int c2gpu_eval_points(void *h, long npts, const double *pts, int nder, int periodic, double *res);

/* Kinetic energy density G(r) = res%gkin of grd (nder >= 1) for wavefunction fields: 1/2 sum_mo occ |grad phi_mo|^2
   (rho2, src/wfn_private@proc.F90:2104-2106,2209; src/fieldmod@proc.f90:776) -- mesh%fill's im_gkin
   (src/meshmod@proc.f90:436-437,459-460).  Three value-only passes with the basis value replaced by its x, y, z derivative.
   gkin: npts doubles.  Promolecular handles return -1 (the reference has no gkin for them: avail_gkin stays false). */
int c2gpu_eval_points_gkin(void *h, long npts, const double *pts, int periodic, double *gkin);
/* MO values: grd(nder = -1, fder) -> calculate_mo (src/wfn_private@proc.F90:2447-2553), mesh%fill's prop > 100
   (src/meshmod@proc.f90:461-464): phi(x) = sum_p cmo(imo,p) chi_p(x) over the primitives inside their dran.
   h_mo is a wavefunction handle made by c2gpu_wfn_create over the same primitives with nmoocc = 1, ldc = 1 and
   cmo = the row cmo(imo,:) (the caller resolves fder -> imo as :2461-2530 do), occ = 1.  phi: npts doubles. */
int c2gpu_eval_points_mo(void *h_mo, long npts, const double *pts, int periodic, double *phi);

/* ---- diagnostics used by the roofline accounting -------------------------------------------- */
/* Sum over the grid slab of n_act(p) = #{primitives with dist <= dran} (the reference's own screen, :2747). */
int c2gpu_wfn_count_active(void *h, const double x0[3], const double xd[9], const int nn[3], int periodic, int iz0,
                           int iz1, double *sum_nact);
/* Sum over the grid of the number of atoms inside the promolecular cutoffs (src/environmod@proc.f90:1284-1300). */
int c2gpu_promol_count_active(void *h, const double x0[3], const double xd[9], const int nn[3], int periodic,
                              double *sum_nat);
/* FP64 peaks of the first device, measured now (~0.1 s): chains of independent mma.sync m8n8k4 f64 (DMMA, the tensor
   roofline of the wavefunction kernels) and of independent DFMAs (the roofline of the primitive / promolecular kernels).
   bench.py prints them beside the recorded figures it normalises with. */
int c2gpu_fp64_peak(double *dmma_tflops, double *dfma_tflops);
/* Kernel configuration chosen for a wavefunction handle: points per tile, MO columns per block, blocks, threads. */
int c2gpu_wfn_config(void *h, int cfg[8]);
/* The configuration id wfn_create picks for an occupied-MO count (no device needed): 0 S1, 7 S3w, 8 M2, 4 L1, 6 WS. */
int c2gpu_wfn_choose_config(int nmoocc);
/* Override tunables (name/value); used by bench sweeps.  Unknown names return 1. */
int c2gpu_set_option(const char *name, int value);

#ifdef __cplusplus
}
#endif
#endif
