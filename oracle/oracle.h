/*
 * oracle.h -- CPU restatement (plain C + OpenMP) of critic2's grid field-evaluation path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it,
 * and only as the checker / timed CPU baseline.  The product path (critic2_b200/) never
 * calls into this library and has no CPU fallback.
 *
 * PARITY STATUS: the reference (Fortran 2008) cannot be compiled in the authoring container
 * (no Fortran compiler) and the reference's own CUBE/NCIPLOT regression scripts are absent
 * from the snapshot, so this oracle is pinned by physical known-answer
 * tests (electron counts 10/42/18 from the shipped fchk files, analytic single-primitive
 * values, finite-difference consistency, promolecular charge = sum Z), not by reference
 * golden vectors: "parity unpinned" against the real binary.
 *
 * All file:line citations are relative to /root/reference/.
 * Arrays are 0-based here; atom / species / primitive ids keep the reference's 1-based values.
 * All 3x3 matrices are column-major (Fortran order): m[i + 3*j] = M(i+1,j+1).
 */
#ifndef C2_ORACLE_H
#define C2_ORACLE_H

#ifdef __cplusplus
extern "C" {
#endif

/* ---- environment: src/environmod.f90:86-151 (type environ) ---- */
typedef struct orc_environ {
  int ismolecule;
  int n;      /* atoms in the environment (cell + periodic images) */
  int ncell;  /* atoms in the unit cell / molecule */
  int nspc;
  int *zspc;  /* nspc atomic numbers */
  double *x;  /* 3*n reduced crystallographic coordinates */
  double *r;  /* 3*n Cartesian coordinates */
  int *is;    /* n species ids (1-based) */
  int *cidx;  /* n complete-list ids (1-based) */
  int *lvec;  /* 3*n lattice vectors */
  double m_xr2c[9], m_c2xr[9], m_x2xr[9], m_xr2x[9], m_x2c[9], m_c2x[9];
  double x0[3];
  double boxsize, dmax0, rsph_uc, rsph_env;
  int nmin[3], nmax[3], nreg[3], nregion;
  int *imap, *nrlo, *nrhi; /* imap 0-based atom index; nrlo/nrhi 1-based positions into imap like the reference */
  int rs_imax, rs_2imax1, rs_nreg;
  int *rs_ioffset;
  double *rs_rcut;
} orc_environ;

orc_environ *orc_environ_build(int ismol, int nspc, const int *zspc, int ncell, const double *xcryst /*3*ncell*/,
                               const int *is, const double *m_x2c, const double *m_x2xr, const double *m_xr2c,
                               double dmax0_in /* <0: derive as the reference does */);
void orc_environ_free(orc_environ *e);
/* list_near_atoms with up2dcidx (mode 0: cut[ncell]) or up2dsp (mode 1: cut[nspc] = upper limits, lower = 0).
   eid (0-based env index) and dist must hold e->n entries. Returns ierr. */
int orc_list_near_atoms(const orc_environ *e, const double *xp, int mode, const double *cut, int *nat, int *eid,
                        double *dist, int *lvec);
int orc_identify_atom(const orc_environ *e, const double *x0, double distmax);

/* ---- wavefunction: src/wfn_private.f90:83-131 (type molwfn) ---- */
typedef struct orc_wfn {
  int npri, nmoocc, nmoall, wfntyp, nalpha, nedf, natom;
  int *icenter, *itype;
  double *e, *occ, *cmo; /* cmo column-major (nmoall x npri) */
  int *icenter_edf, *itype_edf;
  double *e_edf, *c_edf;
  /* complete_struct outputs */
  double *dran, *spcutoff, *dran_edf;
  int *lmax, *icord, *iprilo, *iprihi, *icord_edf, *iprilo_edf, *iprihi_edf; /* 1-based values */
  double globalcutoff;
} orc_wfn;

orc_wfn *orc_wfn_create(int natom, int npri, const int *icenter, const int *itype, const double *e, int nmoocc,
                        int nmoall, const double *cmo, const double *occ, int wfntyp, int nalpha, int nedf,
                        const int *icenter_edf, const int *itype_edf, const double *e_edf, const double *c_edf);
void orc_wfn_free(orc_wfn *f);
/* accessors for the complete_struct tables (tests compare them with the host-side mirror) */
void orc_wfn_tables(const orc_wfn *f, double *dran, double *spcutoff, int *lmax, int *icord, int *iprilo, int *iprihi);

/* rho2: out[0..2]=rho(3) [3..5]=rhoval(3) [6..14]=grad(3,3) [15..23]=gradval(3,3) [24..50]=h(3,3,3)
   [51..77]=hval [78..80]=gkin(3) [81]=vir [82..90]=stress(3,3); all Fortran order. */
void orc_rho2(const orc_wfn *f, const orc_environ *env, const double *xpos, int nder, double *out91);
/* calculate_mo (:2447-2553) with a numeric MO selector (1-based, <= nmoall): the MO value at xpos */
double orc_calculate_mo(const orc_wfn *f, const orc_environ *env, const double *xpos, int imo);
/* number of primitives passing the reference's own per-point screen (:2747) -- the n_act of SURVEY §8(d) */
long orc_nact(const orc_wfn *f, const orc_environ *env, const double *xpos);

/* ---- radial grids: src/grid1mod.f90:25-41 ---- */
typedef struct orc_grid1 {
  int isinit, ngrid, z, qat;
  double a, b, rmax, rmax2;
  double *r, *f, *fp, *fpp;
} orc_grid1;
orc_grid1 *orc_grid1_from_table(int z, int qat, int ngrid, double a, double b, const double *r, const double *f,
                                const double *fp, const double *fpp);
/* read_critic (grid1mod@proc.f90:204-332): build from a dat/wfc/<el>_pbe.wfc file; n = electrons used */
orc_grid1 *orc_grid1_read_critic(const char *file, int z, int q);
void orc_grid1_free(orc_grid1 *g);
void orc_grid1_interp(const orc_grid1 *g, double r0, double *f, double *fp, double *fpp);

/* ---- promolecular: src/environmod@proc.f90:1202-1327 ---- */
/* agrid: array indexed by atomic number 1..118 (entry 0 unused), NULL = not initialised.
   fragmask: NULL or ncell flags (1 = atom's cidx belongs to the fragment). x0 Cartesian. */
void orc_promolecular(const orc_environ *e, orc_grid1 *const *agrid, const double *x0, int nder, const unsigned char *fragmask,
                      double *f, double *fp, double *fpp);

/* ---- field: src/fieldmod@proc.f90:613-928 ---- */
enum { ORC_TYPE_WFN = 0, ORC_TYPE_PROMOL = 1 };
typedef struct orc_field {
  int type;
  const orc_environ *env; /* crystal env (sy%c%env); for wfn also f%wfn%env */
  const orc_wfn *wfn;
  orc_grid1 *const *agrid;
  const unsigned char *fragmask; /* type_promol_frag */
  /* f%usecore with f%zpsp (src/fieldmod.f90:64,79): the core tables cgrid(iz,zpsp(iz)) indexed by atomic number
     (NULL entry: zpsp(iz) <= 0, no core for that element); NULL array = usecore off.  grd :827-832, grd0 :922-925 */
  orc_grid1 *const *cgrid;
} orc_field;
/* scalar_value subset (src/types.f90:113-148): res[0]=f, [1..3]=gf, [4..12]=hf (col-major), [13]=gfmod,
   [14]=del2f, [15]=isnuc, [16]=fval, [17]=gkin, [18]=vir, [19..27]=stress, [28..29]=fspin */
#define ORC_RES_LEN 30
void orc_grd(const orc_field *f, const double *v, int nder, int periodic, double *res);
double orc_grd0(const orc_field *f, const double *v, int periodic);

/* eig -> trace/tredig/tqlgrm (src/tools_math@proc.f90:822, src/oldlibs/{trace,tredig,tqlgrm}.f): h is n x n
   col-major, overwritten by eigenvectors; eval ascending; returns ierr */
int orc_eig(double *h, double *eval, int n);

/* ---- drivers ---- */
/* rhoplot_cube evaluation loop (src/rhoplot@proc.f90:582-631): one array per set bit of outmask (bit nti, nti=0..11
   = f,gx,gy,gz,gmod,hxx,hxy,hxz,hyy,hyz,hzz,lap), each nn1*nn2*nn3, x fastest. */
void orc_cube(const orc_field *f, const double *x0, const double *xd /*3x3 col-major*/, const int *nn, int periodic,
              unsigned outmask, double *const *out, int nthreads);
/* nciplot evaluation loop (src/nci@proc.f90:499-561); arrays are (k,j,i) with k (z) fastest. rhoat may be NULL. */
void orc_nci(const orc_field *fref, const orc_field *fpromol, const double *x0, const double *xmat, const int *nstep,
             double *crho, double *cgrad, double *rhoat, int nthreads);
/* nciplot cutoff pass (src/nci@proc.f90:569-604), without fragments. dat (2*ndat doubles) may be NULL.
   sums[3] = sumq, sumqp, sumv (already scaled by omega/N). Returns the number of .dat pairs. */
long orc_nci_cutoffs(const int *nstep, double *crho, double *cgrad, const double *rhoat, double rhocut, double dimcut,
                     double rhoplot, const double *srhorange, int onlyneg, double rho_void, double omega, double *sums,
                     double *dat, long maxdat);
/* nciplot fragment densities (src/nci@proc.f90:534-550) and the cutoff pass with fragments (:579-582);
   rhofrag is [nfrag][n1*n2*n3] */
void orc_nci_fragments(int nfrag, const int *nat, const double *xyz, const int *z, orc_grid1 *const *agrid,
                       const double *x0, const double *xmat, const int *nstep, double *rhofrag, int nthreads);
long orc_nci_cutoffs_frag(const int *nstep, double *crho, double *cgrad, const double *rhoat, int nfrag,
                          const double *rhofrag, double rhoparam, double rhoparam2, double rhocut, double dimcut,
                          double rhoplot, const double *srhorange, int onlyneg, double rho_void, double omega,
                          double *sums, double *dat, long maxdat);
/* promolecular_grid (src/crystalmod@proc.f90:5118-5158): whole-cell grid at fractional nodes (i-1)/n, nder=0 */
void orc_promolecular_grid(const orc_environ *e, orc_grid1 *const *agrid, const int *n, const unsigned char *fragmask,
                           double *out, int nthreads);
/* number of atoms entering the radial sum of promolecular at each point (the pairs F_pro of SURVEY 8(d) counts) */
void orc_promol_nat_points(const orc_environ *e, orc_grid1 *const *agrid, long npts, const double *pts,
                           const unsigned char *fragmask, long *nat, int nthreads);
/* formatted cube bodies (oracle_text.c): writegrid_cube / write_cube_body / .dat text */
long orc_text_record_bytes(int nval, int fmt);
void orc_cube_text(const double *g, const int *n, int fmt, int order, char *text, int nthreads);
/* evaluate grd on an arbitrary list of points: res is npts*ORC_RES_LEN */
void orc_grd_points(const orc_field *f, long npts, const double *pts, int nder, int periodic, double *res, int nthreads);
/* grd(v, nder=-1, fder=imo) on a point list (src/fieldmod@proc.f90:791-793): wrap like grd, then calculate_mo at wcr */
void orc_mo_points(const orc_field *f, long npts, const double *pts, int imo, int periodic, double *phi, int nthreads);
void orc_nact_points(const orc_wfn *f, const orc_environ *env, long npts, const double *pts, long *nact, int nthreads);

double orc_cutrad(int z);

#ifdef __cplusplus
}
#endif
#endif
