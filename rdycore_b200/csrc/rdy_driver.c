/* rdy_driver.c — RDycore's public lifecycle (include/rdycore_b200.h) driving the B200 SWE operator.  Host C only.
 *
 * RDyCreate  : reads the scalar sections of an RDycore deck (physics.flow, numerics, time)        src/yaml_input.c:40-190, 845-1000
 * RDySetup   : builds the device operator from the mesh handed in by RDyB200SetMesh                src/rdysetup.c:1373-1575
 * RDyAdvance : one coupling interval: adaptive dt from the last Courant number, then TS steps      src/rdyadvance.c:261-383
 *              with TS_EXACTFINALTIME_MATCHSTEP (the last step is shortened to land on the interval end)
 * The stage loop that PETSc TS runs on the host (one RHS callback per stage) runs here as a queue of fused kernels on
 * the operator's stream; the host waits only when a getter or the adaptive-dt rule needs a number.
 */
#include "rdycore_b200.h"

#include <ctype.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "rdyb200_io.h"
#include "rdyb_error.h"

#define INVALID_REAL (-999.0) /* src/yaml_input.c: the "not given" markers */
#define INVALID_INT (-999)
#define MAX_DECK_ENTRIES 256

typedef struct {
  char key[160], value[96];
} DeckEntry;

typedef struct {
  int32_t  num_owned;
  int32_t *owned_cell_ids;
} Region;

struct _p_RDy {
  MPI_Comm comm;
  char     config_file[1024];
  /* config (RDyConfig, include/private/rdyconfigimpl.h) */
  RdybConfig  physics;
  RDyTimeUnit unit;
  double      stop, time_step, coupling_interval; /* in `unit` */
  int32_t     stop_n;
  int         adaptive;
  double      target_courant, max_increase_factor, initial_time_step;
  /* mesh / boundaries (borrowed) */
  RdybHostMesh      mesh;
  int32_t           num_boundaries;
  RdybBoundaryView *boundaries; /* copied descriptors; edge id arrays borrowed */
  int32_t           num_owned, num_local;
  Region           *regions;
  int32_t           num_regions;
  /* device */
  int          device, kernel_kind, tile_cells;
  RdybOperator op;
  int          is_setup;
  /* TS state */
  double  time, dt; /* seconds */
  int32_t step;
  int     courant_valid;
  double  last_courant;
  /* device-side time loop (rdyb_operator_advance_device): when it runs, dt / step / Courant live in device memory and the fields above
   * are refreshed only when a getter asks (time_refresh) */
  int     dev_loop;   /* the last RDyAdvance ran on the device: dt, step, last_courant above are stale */
  int     ctl_stale;  /* the host path (or a restart) changed time / dt: push them to the device before the next device interval */
  int32_t step_base;  /* step count at the last rdyb_operator_time_set */
};

static int g_initialized = 0;

/* ------------------------------------------------------------------ process lifecycle ---- */
PetscErrorCode RDyInit(int argc, char *argv[], const char *help) {
  (void)argc; (void)argv; (void)help;
  g_initialized = 1;
  return PETSC_SUCCESS;
}
PetscInt RDyFinalize(void) {
  g_initialized = 0;
  return 0;
}
PetscBool RDyInitialized(void) { return g_initialized ? PETSC_TRUE : PETSC_FALSE; }

/* ------------------------------------------------------------------ time units (src/rdysetup.c:41-105) ---- */
static double unit_seconds(RDyTimeUnit u) {
  switch (u) {
    case RDY_TIME_SECONDS: return 1.0;
    case RDY_TIME_MINUTES: return 60.0;
    case RDY_TIME_HOURS: return 3600.0;
    case RDY_TIME_DAYS: return 86400.0;
    case RDY_TIME_MONTHS: return 30.0 * 86400.0;
    case RDY_TIME_YEARS: return 365.0 * 86400.0;
    default: return 0.0;
  }
}
static double to_seconds(double t, RDyTimeUnit u) { return t * unit_seconds(u); }
static double from_seconds(double t, RDyTimeUnit u) { return t / unit_seconds(u); }

PetscErrorCode RDyConvertTime(RDyTimeUnit unit_from, PetscReal t_from, RDyTimeUnit unit_to, PetscReal *t_to) {
  RDYB_CHECK(t_to, RDYB_ERR_ARG, "NULL argument");
  RDYB_CHECK(unit_seconds(unit_from) > 0.0 && unit_seconds(unit_to) > 0.0, RDYB_ERR_USER, "Unsupported time unit");
  *t_to = from_seconds(to_seconds(t_from, unit_from), unit_to);
  return PETSC_SUCCESS;
}

/* ------------------------------------------------------------------ the deck: nested maps with scalar leaves ---- */
/* Flattens "a:\n  b:\n    c: v" into ("a.b.c", "v").  List items ("- ...") and everything nested under them are skipped:
 * those sections (regions, materials, conditions, output ...) are out of scope and arrive through calls instead. */
static int deck_read(const char *path, DeckEntry *entries, int *count) {
  FILE *f = fopen(path, "r");
  RDYB_CHECK(f, RDYB_ERR_USER, "Could not open configuration file %s", path); /* src/rdycore.c: PETSC_ERR_FILE_OPEN-like */
  char  line[512], stack[8][64];
  int   indent_of[8], depth = 0, n = 0, skip_indent = -1;
  while (fgets(line, sizeof(line), f)) {
    char *hash = strchr(line, '#');
    if (hash) *hash = '\0';
    int len = (int)strlen(line);
    while (len > 0 && isspace((unsigned char)line[len - 1])) line[--len] = '\0';
    int indent = 0;
    while (line[indent] == ' ') ++indent;
    if (line[indent] == '\0') continue;
    if (skip_indent >= 0) {
      if (indent > skip_indent || line[indent] == '-') continue; /* inside a list section */
      skip_indent = -1;
    }
    while (depth > 0 && indent <= indent_of[depth - 1]) --depth;
    if (line[indent] == '-') { /* a list directly under the current key: skip it */
      skip_indent = depth > 0 ? indent_of[depth - 1] : 0;
      continue;
    }
    char *colon = strchr(line + indent, ':');
    if (!colon) { fclose(f); return rdyb_fail(RDYB_ERR_USER, "%s: cannot parse line '%s'", path, line + indent); }
    *colon = '\0';
    char *key = line + indent, *val = colon + 1;
    int klen = (int)strlen(key);
    while (klen > 0 && isspace((unsigned char)key[klen - 1])) key[--klen] = '\0';
    while (*val == ' ') ++val;
    if (*val == '\0') { /* a map opens */
      if (depth == 8) { fclose(f); return rdyb_fail(RDYB_ERR_USER, "%s: maps nested deeper than 8", path); }
      snprintf(stack[depth], sizeof(stack[depth]), "%s", key);
      indent_of[depth++] = indent;
      continue;
    }
    if (*val == '[' || *val == '{') continue; /* inline collections belong to out-of-scope sections */
    if (n == MAX_DECK_ENTRIES) { fclose(f); return rdyb_fail(RDYB_ERR_USER, "%s: more than %d scalar entries", path, MAX_DECK_ENTRIES); }
    size_t at = 0;
    entries[n].key[0] = '\0';
    for (int d = 0; d < depth && at < sizeof(entries[n].key); ++d) at += (size_t)snprintf(entries[n].key + at, sizeof(entries[n].key) - at, "%s.", stack[d]);
    if (at < sizeof(entries[n].key)) snprintf(entries[n].key + at, sizeof(entries[n].key) - at, "%s", key);
    snprintf(entries[n].value, sizeof(entries[n].value), "%s", val);
    ++n;
  }
  fclose(f);
  *count = n;
  return RDYB_SUCCESS;
}

static const char *deck_get(const DeckEntry *e, int n, const char *key) {
  for (int i = 0; i < n; ++i)
    if (!strcmp(e[i].key, key)) return e[i].value;
  return NULL;
}
static int deck_real(const DeckEntry *e, int n, const char *key, double *out) {
  const char *v = deck_get(e, n, key);
  if (!v) return RDYB_SUCCESS;
  char *end;
  const double x = strtod(v, &end);
  RDYB_CHECK(end != v && *end == '\0', RDYB_ERR_USER, "Invalid value for %s: '%s'", key, v);
  *out = x;
  return RDYB_SUCCESS;
}
static int deck_enum(const DeckEntry *e, int n, const char *key, const char *const *names, const int *values, int count, int *out) {
  const char *v = deck_get(e, n, key);
  if (!v) return RDYB_SUCCESS;
  for (int i = 0; i < count; ++i)
    if (!strcmp(v, names[i])) { *out = values[i]; return RDYB_SUCCESS; }
  return rdyb_fail(RDYB_ERR_USER, "Invalid value for %s: '%s'", key, v);
}

/* ------------------------------------------------------------------ RDyCreate ---- */
static int read_config(RDy rdy) {
  static DeckEntry entries[MAX_DECK_ENTRIES];
  int n = 0;
  RDYB_CALL(deck_read(rdy->config_file, entries, &n));
  /* physics.flow (src/yaml_input.c:40-90) */
  const char *mode = deck_get(entries, n, "physics.flow.mode");
  RDYB_CHECK(!mode || !strcmp(mode, "swe"), RDYB_ERR_USER, "Only the 'swe' flow mode is currently supported."); /* src/rdysetup.c:1200 */
  rdy->physics.tiny_h = 0.0; rdy->physics.h_anuga_regular = 0.0; rdy->physics.xq2018_threshold = 0.0;
  RDYB_CALL(deck_real(entries, n, "physics.flow.tiny_h", &rdy->physics.tiny_h));
  RDYB_CALL(deck_real(entries, n, "physics.flow.h_anuga_reg_parameter", &rdy->physics.h_anuga_regular));
  RDYB_CALL(deck_real(entries, n, "physics.flow.source.xq2018_threshold", &rdy->physics.xq2018_threshold));
  {
    static const char *const names[] = {"semi_implicit", "implicit_xq2018"};
    static const int vals[] = {RDYB_SOURCE_SEMI_IMPLICIT, RDYB_SOURCE_IMPLICIT_XQ2018};
    int v = RDYB_SOURCE_SEMI_IMPLICIT;
    RDYB_CALL(deck_enum(entries, n, "physics.flow.source.method", names, vals, 2, &v));
    rdy->physics.source_method = v;
  }
  {
    static const char *const names[] = {"none", "bradford_sanders_2002", "hydrostatic_reconstruction"};
    static const int vals[] = {RDYB_WELL_BALANCING_NONE, RDYB_WELL_BALANCING_BS2002, RDYB_WELL_BALANCING_HR};
    int v = RDYB_WELL_BALANCING_NONE;
    RDYB_CALL(deck_enum(entries, n, "physics.flow.well_balancing", names, vals, 3, &v));
    /* the PETSc backend runs BS2002 decks with the standard operators (src/operator.c:202-209) */
    rdy->physics.well_balancing = v == RDYB_WELL_BALANCING_BS2002 ? RDYB_WELL_BALANCING_NONE : v;
  }
  /* defaults, src/yaml_input.c:854-865 */
  if (!rdy->physics.tiny_h) rdy->physics.tiny_h = 1e-7;
  if (rdy->physics.xq2018_threshold > 0.0 && rdy->physics.source_method != RDYB_SOURCE_IMPLICIT_XQ2018)
    return rdyb_fail(RDYB_ERR_USER, "Erroneously set XQ2018 threshold: %g", rdy->physics.xq2018_threshold); /* :910-912 */
  if (rdy->physics.source_method == RDYB_SOURCE_IMPLICIT_XQ2018 && !rdy->physics.xq2018_threshold) rdy->physics.xq2018_threshold = 1e-10;
  /* numerics (src/yaml_input.c:100-150) */
  {
    static const char *const sp[] = {"fv"};
    static const int spv[] = {0};
    int v = 0;
    RDYB_CALL(deck_enum(entries, n, "numerics.spatial", sp, spv, 1, &v));
    static const char *const tn[] = {"euler", "rk4"};
    static const int tv[] = {RDYB_TEMPORAL_EULER, RDYB_TEMPORAL_RK4};
    v = RDYB_TEMPORAL_EULER;
    const char *t = deck_get(entries, n, "numerics.temporal");
    RDYB_CHECK(!t || strcmp(t, "beuler"), RDYB_ERR_SUP, "numerics.temporal: beuler (implicit, SNES) is outside the explicit B200 path");
    RDYB_CALL(deck_enum(entries, n, "numerics.temporal", tn, tv, 2, &v));
    rdy->physics.temporal = v;
    /* numerics.second_order / limiter / no_limiter (src/yaml_input.c:130-143; resolved as src/swe/swe_petsc.c:357-369) */
    static const char *const bn[] = {"true", "false"};
    static const int bv[] = {1, 0};
    int second = 0, nolim = 0, lim = 0;
    RDYB_CALL(deck_enum(entries, n, "numerics.second_order", bn, bv, 2, &second));
    RDYB_CALL(deck_enum(entries, n, "numerics.no_limiter", bn, bv, 2, &nolim));
    static const char *const ln[] = {"minmod", "none", "van_leer"};
    static const int lv[] = {0, 1, 2};
    RDYB_CALL(deck_enum(entries, n, "numerics.limiter", ln, lv, 3, &lim));
    if (nolim) lim = 1;
    rdy->physics.reconstruction = !second ? RDYB_RECON_FIRST_ORDER : lim == 1 ? RDYB_RECON_MUSCL_NONE : lim == 2 ? RDYB_RECON_MUSCL_VANLEER : RDYB_RECON_MUSCL_MINMOD;
    static const char *const rn[] = {"roe"};
    static const int rv[] = {0};
    RDYB_CALL(deck_enum(entries, n, "numerics.riemann", rn, rv, 1, &v));
  }
  /* time (src/yaml_input.c:150-190, 955-1000) */
  {
    static const char *const un[] = {"seconds", "minutes", "hours", "days", "months", "years"};
    static const int uv[] = {RDY_TIME_SECONDS, RDY_TIME_MINUTES, RDY_TIME_HOURS, RDY_TIME_DAYS, RDY_TIME_MONTHS, RDY_TIME_YEARS};
    int u = RDY_TIME_UNSET;
    RDYB_CALL(deck_enum(entries, n, "time.unit", un, uv, 6, &u));
    RDYB_CHECK(u != RDY_TIME_UNSET, RDYB_ERR_USER, "time.unit must be specified");
    rdy->unit = (RDyTimeUnit)u;
  }
  rdy->stop = INVALID_REAL; rdy->time_step = INVALID_REAL; rdy->coupling_interval = INVALID_REAL; rdy->stop_n = INVALID_INT;
  RDYB_CALL(deck_real(entries, n, "time.stop", &rdy->stop));
  RDYB_CALL(deck_real(entries, n, "time.time_step", &rdy->time_step));
  RDYB_CALL(deck_real(entries, n, "time.coupling_interval", &rdy->coupling_interval));
  {
    double sn = INVALID_INT;
    RDYB_CALL(deck_real(entries, n, "time.stop_n", &sn));
    rdy->stop_n = (int32_t)sn;
  }
  {
    static const char *const bn[] = {"true", "false"};
    static const int bv[] = {1, 0};
    RDYB_CALL(deck_enum(entries, n, "time.adaptive.enable", bn, bv, 2, &rdy->adaptive));
  }
  if (rdy->adaptive) {
    rdy->target_courant = 0.0; rdy->max_increase_factor = 0.0; rdy->initial_time_step = 0.0;
    RDYB_CALL(deck_real(entries, n, "time.adaptive.target_courant_number", &rdy->target_courant));
    RDYB_CALL(deck_real(entries, n, "time.adaptive.max_increase_factor", &rdy->max_increase_factor));
    RDYB_CALL(deck_real(entries, n, "time.adaptive.initial_time_step", &rdy->initial_time_step));
    RDYB_CHECK(rdy->target_courant > 0.0, RDYB_ERR_USER, "time.adaptive.target_courant_number must be greater than 0.0");
    RDYB_CHECK(rdy->target_courant < 1.0, RDYB_ERR_USER, "time.adaptive.target_courant_number must be less than 1.0");
    RDYB_CHECK(rdy->max_increase_factor > 1.0, RDYB_ERR_USER, "time.adaptive.max_increase_factor must be greater than 1.0");
    RDYB_CHECK(rdy->initial_time_step > 0.0, RDYB_ERR_USER, "time.adaptive.initial_time_step must be greater than 0.0");
    RDYB_CHECK(rdy->stop_n == INVALID_INT, RDYB_ERR_USER, "max_step cannot be specified with adaptive time step enabled");
    RDYB_CHECK(rdy->time_step == INVALID_REAL, RDYB_ERR_USER, "time_step cannot be specified with adaptive time step enabled");
    rdy->time_step = rdy->initial_time_step;
  }
  int given = (rdy->stop != INVALID_REAL) + (rdy->stop_n != INVALID_INT) + (rdy->time_step != INVALID_REAL);
  RDYB_CHECK(given == 2, RDYB_ERR_USER, "Exactly 2 of time.final_time, time.max_step, time.time_step must be specified (%d given)", given);
  if (rdy->stop == INVALID_REAL) rdy->stop = rdy->stop_n * rdy->time_step;
  else if (rdy->stop_n == INVALID_INT) rdy->stop_n = (int32_t)(rdy->stop / rdy->time_step);
  else rdy->time_step = rdy->stop / rdy->stop_n;
  if (rdy->coupling_interval == INVALID_REAL) rdy->coupling_interval = rdy->stop;
  else {
    RDYB_CHECK(rdy->coupling_interval > 0.0, RDYB_ERR_USER, "time.coupling_interval must be positive");
    RDYB_CHECK(rdy->coupling_interval <= rdy->stop, RDYB_ERR_USER, "time.coupling_interval must not exceed time.final_time");
  }
  return RDYB_SUCCESS;
}

PetscErrorCode RDyCreate(MPI_Comm comm, const char *config_file, RDy *rdy) {
  RDYB_CHECK(rdy && config_file, RDYB_ERR_ARG, "RDyCreate: NULL argument");
  RDYB_CHECK(g_initialized, RDYB_ERR_USER, "RDyInit must be called before RDyCreate");
  RDy r = calloc(1, sizeof(*r));
  RDYB_CHECK(r, RDYB_ERR_MEM, "out of memory");
  r->comm = comm;
  snprintf(r->config_file, sizeof(r->config_file), "%s", config_file);
  r->device = 0; r->kernel_kind = 3; r->tile_cells = 224;
  int ierr = read_config(r);
  if (ierr) { free(r); return ierr; }
  *rdy = r;
  return PETSC_SUCCESS;
}

PetscErrorCode RDyDestroy(RDy *rdy) {
  if (!rdy || !*rdy) return PETSC_SUCCESS;
  RDy r = *rdy;
  int ierr = rdyb_operator_destroy(&r->op);
  for (int32_t i = 0; i < r->num_regions; ++i) free(r->regions[i].owned_cell_ids);
  free(r->regions);
  free(r->boundaries);
  free(r);
  *rdy = NULL;
  return ierr;
}

/* ------------------------------------------------------------------ mesh, regions, device ---- */
PetscErrorCode RDyB200SetMesh(RDy rdy, RdybHostMesh mesh, PetscInt nb, const RdybBoundaryView *boundaries) {
  RDYB_CHECK(rdy && mesh, RDYB_ERR_ARG, "RDyB200SetMesh: NULL argument");
  RDYB_CHECK(!rdy->is_setup, RDYB_ERR_USER, "RDyB200SetMesh must be called before RDySetup");
  RDYB_CHECK(nb >= 0 && (nb == 0 || boundaries), RDYB_ERR_ARG, "RDyB200SetMesh: %d boundaries but no boundary array", nb);
  int32_t sizes[6];
  RDYB_CALL(rdyb_hostmesh_sizes(mesh, sizes));
  rdy->mesh = mesh; rdy->num_local = sizes[1]; rdy->num_owned = sizes[2];
  free(rdy->boundaries);
  rdy->boundaries = calloc((size_t)nb + 1, sizeof(RdybBoundaryView));
  RDYB_CHECK(rdy->boundaries, RDYB_ERR_MEM, "out of memory");
  for (PetscInt b = 0; b < nb; ++b) rdy->boundaries[b] = boundaries[b];
  rdy->num_boundaries = nb;
  return PETSC_SUCCESS;
}

PetscErrorCode RDyB200AddRegion(RDy rdy, PetscInt n, const PetscInt *ids, PetscInt *region_idx) {
  RDYB_CHECK(rdy && (n == 0 || ids), RDYB_ERR_ARG, "RDyB200AddRegion: NULL argument");
  Region *r = realloc(rdy->regions, sizeof(Region) * (size_t)(rdy->num_regions + 1));
  RDYB_CHECK(r, RDYB_ERR_MEM, "out of memory");
  rdy->regions = r;
  Region *g = &r[rdy->num_regions];
  g->num_owned = n;
  g->owned_cell_ids = malloc(sizeof(int32_t) * (size_t)(n > 0 ? n : 1));
  RDYB_CHECK(g->owned_cell_ids, RDYB_ERR_MEM, "out of memory");
  if (n) memcpy(g->owned_cell_ids, ids, sizeof(int32_t) * (size_t)n);
  if (region_idx) *region_idx = rdy->num_regions;
  ++rdy->num_regions;
  return PETSC_SUCCESS;
}

PetscErrorCode RDyB200SetDevice(RDy rdy, int device) {
  RDYB_CHECK(rdy && !rdy->is_setup, RDYB_ERR_USER, "RDyB200SetDevice must be called before RDySetup");
  rdy->device = device;
  return PETSC_SUCCESS;
}
PetscErrorCode RDyB200SetKernel(RDy rdy, PetscInt kind, PetscInt tile_cells) {
  RDYB_CHECK(rdy, RDYB_ERR_ARG, "NULL RDy");
  rdy->kernel_kind = kind; rdy->tile_cells = tile_cells;
  if (rdy->is_setup) RDYB_CALL(rdyb_operator_set_kernel(rdy->op, kind, tile_cells));
  return PETSC_SUCCESS;
}
PetscErrorCode RDyB200GetOperator(RDy rdy, RdybOperator *op) {
  RDYB_CHECK(rdy && op, RDYB_ERR_ARG, "NULL argument");
  RDYB_CHECK(rdy->is_setup, RDYB_ERR_USER, "RDySetup has not been called");
  *op = rdy->op;
  return PETSC_SUCCESS;
}

/* ------------------------------------------------------------------ RDySetup (src/rdysetup.c:1373-1575) ---- */
PetscErrorCode RDySetup(RDy rdy) {
  RDYB_CHECK(rdy, RDYB_ERR_ARG, "NULL RDy");
  RDYB_CHECK(!rdy->is_setup, RDYB_ERR_USER, "RDySetup was already called");
  RDYB_CHECK(rdy->mesh, RDYB_ERR_USER, "no mesh: call RDyB200SetMesh before RDySetup (mesh files are read by the caller)");
  RdybMeshView view;
  RdybHaloView halo;
  RDYB_CALL(rdyb_hostmesh_view(rdy->mesh, &view));
  if (rdy->physics.well_balancing == RDYB_WELL_BALANCING_HR && !view.cell_zc) {
    /* src/rdysetup.c:1507-1509 + src/swe/swe_petsc.c:1208-1224: 2-D projected geometry, bed elevation = mean vertex z
     * (a caller that read a cell_elevation file has set zc itself) */
    RDYB_CALL(rdyb_hostmesh_override_2d_projection(rdy->mesh));
    RDYB_CALL(rdyb_hostmesh_set_zc(rdy->mesh, NULL));
    RDYB_CALL(rdyb_hostmesh_view(rdy->mesh, &view));
  }
  RDYB_CALL(rdyb_hostmesh_halo(rdy->mesh, &halo));
  RDYB_CALL(rdyb_operator_create(&rdy->physics, &view, rdy->num_boundaries, rdy->boundaries, halo.num_peers > 0 ? &halo : NULL, rdy->device, NULL, &rdy->op));
  /* numerics.second_order runs the tiled MUSCL kernel on 192-cell tiles, the shape K4M is compiled for (swe_tiled_muscl_fixed.cuh), unless RDyB200SetKernel chose a shape */
  if (rdy->physics.reconstruction && rdy->kernel_kind == 3 && rdy->tile_cells == 224) rdy->tile_cells = 192; /* K4M 0.639 ms; K3M: 0.876 ms against 0.987 (128) and 1.089 (224), profiles/r2_so_tile*.json */
  RDYB_CALL(rdyb_operator_set_kernel(rdy->op, rdy->kernel_kind, rdy->tile_cells));
  RDYB_CALL(rdyb_operator_set_options(rdy->op, 1, 0, 0, 0));
  rdy->time = 0.0; rdy->step = 0;
  rdy->dt = to_seconds(rdy->time_step, rdy->unit); /* InitSolver: TSSetTimeStep(config.time.time_step) */
  rdy->courant_valid = 0;
  rdy->dev_loop = 0; rdy->ctl_stale = 1; rdy->step_base = 0;
  rdy->is_setup = 1;
  return PETSC_SUCCESS;
}

/* ------------------------------------------------------------------ RDyAdvance (src/rdyadvance.c:261-383) ---- */
/* dt, step count and the Courant number of the device-side loop -> the host fields (the one place that waits for the device) */
static int time_refresh(RDy rdy) {
  if (!rdy->dev_loop) return RDYB_SUCCESS;
  double  t, dt, courant;
  int64_t steps;
  RDYB_CALL(rdyb_operator_time_get(rdy->op, &t, &dt, &steps, &courant));
  rdy->dt = dt;
  rdy->step = rdy->step_base + (int32_t)steps;
  rdy->last_courant = courant;
  rdy->courant_valid = courant > 0.0;
  (void)t; /* == rdy->time: every interval ends exactly on the coupling time */
  rdy->dev_loop = 0;
  return RDYB_SUCCESS;
}

static int run_steps(RDy rdy, double dt, int32_t nsteps) {
  if (nsteps <= 0) return RDYB_SUCCESS;
  if (rdy->physics.temporal == RDYB_TEMPORAL_RK4) return rdyb_operator_step_rk4(rdy->op, dt, nsteps);
  return rdyb_operator_step_euler(rdy->op, dt, nsteps);
}

PetscErrorCode RDyAdvance(RDy rdy) {
  RDYB_CHECK(rdy, RDYB_ERR_ARG, "NULL RDy");
  RDYB_CHECK(rdy->is_setup, RDYB_ERR_USER, "RDySetup has not been called");
  const double interval = to_seconds(rdy->coupling_interval, rdy->unit);
  const double t_end = rdy->time + interval;
  if (rdy->physics.temporal == RDYB_TEMPORAL_EULER) {
    /* the whole interval - dt rule, step count, last-step shortening, Courant reduction - as ONE graph launch with a device-side
     * loop: no host synchronisation per RDyAdvance.  RDYB_ERR_SUP: this configuration has no device loop (partitioned mesh,
     * second order, a kernel other than K1 / K4); fall through to the host-driven path below. */
    if (rdy->ctl_stale) {
      RDYB_CALL(time_refresh(rdy));
      RDYB_CALL(rdyb_operator_time_set(rdy->op, rdy->time, rdy->dt));
      rdy->step_base = rdy->step; rdy->ctl_stale = 0;
    }
    const int ierr = rdyb_operator_advance_device(rdy->op, interval, rdy->adaptive, rdy->target_courant, rdy->max_increase_factor);
    if (ierr == RDYB_SUCCESS) { rdy->time = t_end; rdy->dev_loop = 1; return PETSC_SUCCESS; }
    if (ierr != RDYB_ERR_SUP) return ierr;
  }
  RDYB_CALL(time_refresh(rdy));
  rdy->ctl_stale = 1;
  if (rdy->adaptive && rdy->courant_valid) { /* :303-343 */
    double dt = rdy->dt;
    if (rdy->last_courant < rdy->target_courant) {
      const double factor = fmin(rdy->target_courant / rdy->last_courant, rdy->max_increase_factor);
      dt *= factor;
      if (dt > interval) dt = interval;
    } else {
      dt *= rdy->target_courant / rdy->last_courant;
    }
    rdy->dt = dt;
  }
  /* TSSolve to t_end with TS_EXACTFINALTIME_MATCHSTEP: whole steps of dt, then one shortened step if a remainder is left.
   * (PETSc's step-size bookkeeping is third-party arithmetic; "parity unpinned" at this boundary) */
  const double dt = rdy->dt, span = t_end - rdy->time;
  int32_t nfull = (int32_t)floor(span / dt * (1.0 + 1e-14) + 1e-12);
  if (nfull < 0) nfull = 0;
  double rest = span - nfull * dt;
  if (rest < 1e-12 * fmax(span, dt)) rest = 0.0;
  /* the Courant number RDycore reports is that of the LAST right-hand-side evaluation (diagnostics are reset in every
   * OperatorRHSFunction call, src/rdysetup.c:1136): reset the device slots just before the last step */
  const int32_t before_last = rest > 0.0 ? nfull : nfull - 1;
  RDYB_CALL(run_steps(rdy, dt, before_last));
  RDYB_CALL(rdyb_operator_reset_diagnostics(rdy->op));
  if (rest > 0.0) RDYB_CALL(run_steps(rdy, rest, 1));
  else if (nfull > 0) RDYB_CALL(run_steps(rdy, dt, 1));
  rdy->step += nfull + (rest > 0.0 ? 1 : 0);
  rdy->time = t_end;
  if (rdy->adaptive) { /* UpdateOperatorDiagnostics, :366-372: the one host sync of the interval */
    RdybCourantDiagnostics d;
    RDYB_CALL(rdyb_operator_get_courant_global(rdy->op, &d)); /* all ranks take the same dt (MPI_Allreduce in the reference) */
    rdy->last_courant = d.max_courant_num;
    rdy->courant_valid = d.max_courant_num > 0.0;
  }
  return PETSC_SUCCESS;
}

PetscBool RDyFinished(RDy rdy) { /* :388-400 */
  if (!rdy) return PETSC_TRUE;
  return (rdy->stop - from_seconds(rdy->time, rdy->unit) < 1e-13) ? PETSC_TRUE : PETSC_FALSE;
}

PetscErrorCode RDyB200GetMaxCourantNumber(RDy rdy, PetscReal *courant) {
  RDYB_CHECK(rdy && courant, RDYB_ERR_ARG, "NULL argument");
  RDYB_CHECK(rdy->is_setup, RDYB_ERR_USER, "RDySetup has not been called");
  RdybCourantDiagnostics d;
  RDYB_CALL(rdyb_operator_get_courant(rdy->op, &d));
  *courant = d.max_courant_num;
  return PETSC_SUCCESS;
}

/* ------------------------------------------------------------------ checkpoints (src/checkpoint.c) ---- */
static int owned_natural_ids(RDy rdy, int64_t **ids) {
  const void *g, *o2l;
  int64_t     ng, no;
  RDYB_CALL(rdyb_hostmesh_array(rdy->mesh, "cell_global_ids", &g, &ng));
  RDYB_CALL(rdyb_hostmesh_array(rdy->mesh, "owned_to_local", &o2l, &no));
  *ids = (int64_t *)malloc(sizeof(int64_t) * (size_t)(rdy->num_owned > 0 ? rdy->num_owned : 1));
  RDYB_CHECK(*ids, RDYB_ERR_MEM, "out of memory");
  for (int32_t i = 0; i < rdy->num_owned; ++i) (*ids)[i] = ((const int64_t *)g)[((const int32_t *)o2l)[i]];
  return PETSC_SUCCESS;
}

PetscErrorCode RDyB200WriteCheckpoint(RDy rdy, const char *filename_base, PetscInt nproc, PetscInt rank, PetscInt num_global_cells) {
  RDYB_CHECK(rdy && filename_base, RDYB_ERR_ARG, "NULL argument");
  RDYB_CHECK(rdy->is_setup, RDYB_ERR_USER, "RDySetup has not been called");
  RDYB_CHECK(nproc >= 1 && rank >= 0 && rank < nproc && num_global_cells >= rdy->num_owned, RDYB_ERR_ARG, "bad partition description");
  char path[4096]; /* PETSC_MAX_PATH_LEN */
  RDYB_CHECK(snprintf(path, sizeof path, "%s.bin", filename_base) < (int)sizeof path, RDYB_ERR_ARG_SIZ, "file name too long");
  double  *u   = (double *)malloc(sizeof(double) * 3 * (size_t)(rdy->num_owned > 0 ? rdy->num_owned : 1));
  int64_t *ids = NULL;
  RDYB_CHECK(u, RDYB_ERR_MEM, "out of memory");
  int ierr = rdyb_operator_get_state(rdy->op, u);
  if (!ierr) ierr = time_refresh(rdy);
  if (!ierr) ierr = owned_natural_ids(rdy, &ids);
  const RdybCheckpointMetadata md = {.nproc = nproc, .t = rdy->time, .dt = rdy->dt, .step = rdy->step};
  if (!ierr) ierr = rdyb_checkpoint_write(path, &md, num_global_cells, 3, (int32_t)sizeof(PetscInt), rdy->num_owned, ids, u, rank == 0);
  free(u);
  free(ids);
  return ierr;
}

PetscErrorCode RDyB200ReadCheckpoint(RDy rdy, const char *filename, PetscBool reinitialize) {
  RDYB_CHECK(rdy && filename, RDYB_ERR_ARG, "NULL argument");
  RDYB_CHECK(rdy->is_setup, RDYB_ERR_USER, "RDySetup has not been called");
  RdybCheckpointMetadata md;
  int64_t                n = 0, *ids = NULL;
  double                *nat = NULL;
  RDYB_CALL(rdyb_checkpoint_read(filename, &md, &n, &nat, NULL));
  int     ierr = owned_natural_ids(rdy, &ids);
  double *u    = (double *)malloc(sizeof(double) * 3 * (size_t)(rdy->num_owned > 0 ? rdy->num_owned : 1));
  if (!ierr && !u) ierr = rdyb_fail(RDYB_ERR_MEM, "out of memory");
  for (int32_t i = 0; !ierr && i < rdy->num_owned; ++i) { /* DMPlexNaturalToGlobal: the rows of the owned cells */
    if (ids[i] < 0 || 3 * ids[i] + 2 >= n) ierr = rdyb_fail(RDYB_ERR_ARG_SIZ, "checkpoint %s has %lld values; cell %lld is not in it", filename, (long long)n, (long long)ids[i]);
    else memcpy(u + 3 * (size_t)i, nat + 3 * (size_t)ids[i], 3 * sizeof(double));
  }
  if (!ierr) ierr = rdyb_operator_set_state(rdy->op, u);
  if (!ierr) { /* ConsumeMetadata */
    rdy->dt = md.dt;
    rdy->time = reinitialize ? 0.0 : md.t;
    rdy->step = reinitialize ? 0 : md.step;
    rdy->courant_valid = 0;
    rdy->dev_loop = 0; rdy->ctl_stale = 1;
  }
  free(u);
  free(ids);
  rdyb_free(nat);
  return ierr;
}

/* ------------------------------------------------------------------ time queries (src/rdyadvance.c:402-456) ---- */
PetscErrorCode RDyGetTimeUnit(RDy rdy, RDyTimeUnit *unit) {
  RDYB_CHECK(rdy && unit, RDYB_ERR_ARG, "NULL argument");
  *unit = rdy->unit;
  return PETSC_SUCCESS;
}
PetscErrorCode RDyGetTime(RDy rdy, RDyTimeUnit unit, PetscReal *time) {
  RDYB_CHECK(rdy && time, RDYB_ERR_ARG, "NULL argument");
  *time = from_seconds(rdy->time, unit);
  return PETSC_SUCCESS;
}
PetscErrorCode RDyGetTimeStep(RDy rdy, RDyTimeUnit unit, PetscReal *dt) {
  RDYB_CHECK(rdy && dt, RDYB_ERR_ARG, "NULL argument");
  RDYB_CALL(time_refresh(rdy));
  *dt = from_seconds(rdy->dt, unit);
  return PETSC_SUCCESS;
}
PetscErrorCode RDyGetStep(RDy rdy, PetscInt *step) {
  RDYB_CHECK(rdy && step, RDYB_ERR_ARG, "NULL argument");
  RDYB_CALL(time_refresh(rdy));
  *step = rdy->step;
  return PETSC_SUCCESS;
}
PetscErrorCode RDyGetCouplingInterval(RDy rdy, RDyTimeUnit unit, PetscReal *interval) {
  RDYB_CHECK(rdy && interval, RDYB_ERR_ARG, "NULL argument");
  return RDyConvertTime(rdy->unit, rdy->coupling_interval, unit, interval);
}
PetscErrorCode RDySetCouplingInterval(RDy rdy, RDyTimeUnit unit, PetscReal interval) {
  RDYB_CHECK(rdy, RDYB_ERR_ARG, "NULL RDy");
  return RDyConvertTime(unit, interval, rdy->unit, &rdy->coupling_interval);
}

/* ------------------------------------------------------------------ data access (src/rdydata.c) ---- */
static int check_boundary_index(RDy rdy, PetscInt b) { /* :24-32 */
  RDYB_CHECK(b < rdy->num_boundaries, RDYB_ERR_USER, "Boundary condition index (%d) exceeds the max number of boundary conditions (%d)", b,
             rdy->num_boundaries);
  RDYB_CHECK(b >= 0, RDYB_ERR_USER, "Boundary condition index (%d) cannot be less than zero.", b);
  return RDYB_SUCCESS;
}
static int check_region_index(RDy rdy, PetscInt r) { /* :46-52 */
  RDYB_CHECK(r < rdy->num_regions, RDYB_ERR_USER, "Region index (%d) exceeds the max number of regions (%d)", r, rdy->num_regions);
  RDYB_CHECK(r >= 0, RDYB_ERR_USER, "Region index (%d) cannot be less than zero.", r);
  return RDYB_SUCCESS;
}
static int check_setup(RDy rdy) {
  RDYB_CHECK(rdy, RDYB_ERR_ARG, "NULL RDy");
  RDYB_CHECK(rdy->is_setup, RDYB_ERR_USER, "RDySetup has not been called");
  return RDYB_SUCCESS;
}

PetscErrorCode RDyGetNumGlobalCells(RDy rdy, PetscInt *n) {
  RDYB_CHECK(rdy && n && rdy->mesh, RDYB_ERR_ARG, "NULL argument");
  *n = rdy->num_owned; /* one rank: global = owned; a partitioned caller sums over ranks */
  return PETSC_SUCCESS;
}
PetscErrorCode RDyGetNumOwnedCells(RDy rdy, PetscInt *n) {
  RDYB_CHECK(rdy && n && rdy->mesh, RDYB_ERR_ARG, "NULL argument");
  *n = rdy->num_owned;
  return PETSC_SUCCESS;
}
PetscErrorCode RDyGetNumBoundaryConditions(RDy rdy, PetscInt *n) {
  RDYB_CHECK(rdy && n, RDYB_ERR_ARG, "NULL argument");
  *n = rdy->num_boundaries;
  return PETSC_SUCCESS;
}
PetscErrorCode RDyGetNumBoundaryEdges(RDy rdy, const PetscInt b, PetscInt *n) {
  RDYB_CHECK(rdy && n, RDYB_ERR_ARG, "NULL argument");
  RDYB_CALL(check_boundary_index(rdy, b));
  *n = rdy->boundaries[b].num_edges;
  return PETSC_SUCCESS;
}
PetscErrorCode RDyGetBoundaryConditionFlowType(RDy rdy, const PetscInt b, PetscInt *bc_type) {
  RDYB_CHECK(rdy && bc_type, RDYB_ERR_ARG, "NULL argument");
  RDYB_CALL(check_boundary_index(rdy, b));
  *bc_type = rdy->boundaries[b].bc_type;
  return PETSC_SUCCESS;
}
PetscErrorCode RDyGetBoundaryID(RDy rdy, const PetscInt b, PetscInt *id) {
  RDYB_CHECK(rdy && id, RDYB_ERR_ARG, "NULL argument");
  RDYB_CALL(check_boundary_index(rdy, b));
  *id = rdy->boundaries[b].id;
  return PETSC_SUCCESS;
}

static int get_state_component(RDy rdy, int c, PetscInt size, PetscReal *values) {
  RDYB_CALL(check_setup(rdy));
  RDYB_CHECK(size == rdy->num_owned, RDYB_ERR_ARG_SIZ, "The size of array is not equal to the number of owned cells"); /* :54-58 */
  return rdyb_operator_get_state_component(rdy->op, c, size, values);
}
PetscErrorCode RDyGetOwnedCellHeights(RDy rdy, const PetscInt size, PetscReal *values) { return get_state_component(rdy, 0, size, values); }
PetscErrorCode RDyGetOwnedCellXMomenta(RDy rdy, const PetscInt size, PetscReal *values) { return get_state_component(rdy, 1, size, values); }
PetscErrorCode RDyGetOwnedCellYMomenta(RDy rdy, const PetscInt size, PetscReal *values) { return get_state_component(rdy, 2, size, values); }

static int get_owned_mesh_array(RDy rdy, const char *name, PetscInt size, PetscReal *values) {
  RDYB_CHECK(rdy && rdy->mesh && values, RDYB_ERR_ARG, "NULL argument");
  RDYB_CHECK(size == rdy->num_owned, RDYB_ERR_ARG_SIZ, "The size of array is not equal to the number of owned cells");
  const void *a, *o2l;
  int64_t     na, no;
  RDYB_CALL(rdyb_hostmesh_array(rdy->mesh, name, &a, &na));
  RDYB_CALL(rdyb_hostmesh_array(rdy->mesh, "owned_to_local", &o2l, &no));
  for (PetscInt i = 0; i < size; ++i) values[i] = ((const double *)a)[((const int32_t *)o2l)[i]];
  return PETSC_SUCCESS;
}
PetscErrorCode RDyGetOwnedCellXCentroids(RDy rdy, const PetscInt size, PetscReal *values) { return get_owned_mesh_array(rdy, "cell_xc", size, values); }
PetscErrorCode RDyGetOwnedCellYCentroids(RDy rdy, const PetscInt size, PetscReal *values) { return get_owned_mesh_array(rdy, "cell_yc", size, values); }
PetscErrorCode RDyGetOwnedCellAreas(RDy rdy, const PetscInt size, PetscReal *values) {
  RDYB_CHECK(rdy && rdy->mesh && values, RDYB_ERR_ARG, "NULL argument");
  RDYB_CHECK(size == rdy->num_owned, RDYB_ERR_ARG_SIZ, "The size of array is not equal to the number of owned cells");
  RdybMeshView view;
  RDYB_CALL(rdyb_hostmesh_view(rdy->mesh, &view));
  for (PetscInt c = 0; c < view.num_cells; ++c)
    if (view.cell_is_owned[c]) values[view.cell_local_to_owned[c]] = view.cell_areas[c];
  return PETSC_SUCCESS;
}

PetscErrorCode RDySetFlowDirichletBoundaryValues(RDy rdy, const PetscInt b, const PetscInt num_edges, const PetscInt ndof, PetscReal *values) { /* :88-106 */
  RDYB_CALL(check_setup(rdy));
  RDYB_CALL(check_boundary_index(rdy, b));
  RDYB_CHECK(rdy->boundaries[b].num_edges == num_edges, RDYB_ERR_USER, "The given number of edges (%d) for boundary with index %d is incorrect (should be %d)",
             num_edges, b, rdy->boundaries[b].num_edges);
  RDYB_CHECK(ndof == 3, RDYB_ERR_USER, "The number of DOFs (%d) for the boundary condition need to be three.", ndof);
  RDYB_CHECK(rdy->boundaries[b].bc_type == RDYB_CONDITION_DIRICHLET, RDYB_ERR_USER,
             "Trying to set dirichlet values for boundary with index %d, but it has a different type (%u)", b, (unsigned)rdy->boundaries[b].bc_type);
  return rdyb_operator_set_boundary_values(rdy->op, b, 0, ndof, num_edges, values);
}

static int set_domain_source(RDy rdy, int c, PetscInt size, PetscReal *values) { /* :325-340 */
  RDYB_CALL(check_setup(rdy));
  RDYB_CHECK(size == rdy->num_owned, RDYB_ERR_USER, "Wrong size (%d) for owned cells in domain (%d)", size, rdy->num_owned);
  return rdyb_operator_set_external_source(rdy->op, c, size, values);
}
PetscErrorCode RDySetDomainWaterSource(RDy rdy, PetscInt size, PetscReal *values) { return set_domain_source(rdy, 0, size, values); }
PetscErrorCode RDySetDomainXMomentumSource(RDy rdy, PetscInt size, PetscReal *values) { return set_domain_source(rdy, 1, size, values); }
PetscErrorCode RDySetDomainYMomentumSource(RDy rdy, PetscInt size, PetscReal *values) { return set_domain_source(rdy, 2, size, values); }

static int set_regional_source(RDy rdy, PetscInt r, int c, PetscInt size, PetscReal *values) { /* :225-250 */
  RDYB_CALL(check_setup(rdy));
  RDYB_CALL(check_region_index(rdy, r));
  const Region *g = &rdy->regions[r];
  RDYB_CHECK(size == g->num_owned, RDYB_ERR_USER, "Wrong size (%d) for owned cells in region %d (%d)", size, r, g->num_owned);
  return rdyb_operator_set_regional_external_source(rdy->op, c, size, g->owned_cell_ids, values);
}
PetscErrorCode RDySetRegionalWaterSource(RDy rdy, const PetscInt r, PetscInt size, PetscReal *values) { return set_regional_source(rdy, r, 0, size, values); }
PetscErrorCode RDySetRegionalXMomentumSource(RDy rdy, const PetscInt r, PetscInt size, PetscReal *values) { return set_regional_source(rdy, r, 1, size, values); }
PetscErrorCode RDySetRegionalYMomentumSource(RDy rdy, const PetscInt r, PetscInt size, PetscReal *values) { return set_regional_source(rdy, r, 2, size, values); }

PetscErrorCode RDySetDomainManningsN(RDy rdy, const PetscInt size, PetscReal *values) { /* :524-540 */
  RDYB_CALL(check_setup(rdy));
  RDYB_CHECK(size == rdy->num_owned, RDYB_ERR_USER, "Wrong size (%d) for owned cells in domain (%d)", size, rdy->num_owned);
  return rdyb_operator_set_mannings(rdy->op, size, values);
}
PetscErrorCode RDySetRegionalManningsN(RDy rdy, const PetscInt r, const PetscInt size, PetscReal *values) { /* :505-520 */
  RDYB_CALL(check_setup(rdy));
  RDYB_CALL(check_region_index(rdy, r));
  const Region *g = &rdy->regions[r];
  RDYB_CHECK(size == g->num_owned, RDYB_ERR_USER, "Wrong size (%d) for owned cells in region %d (%d)", size, r, g->num_owned);
  return rdyb_operator_set_regional_mannings(rdy->op, size, g->owned_cell_ids, values);
}

PetscErrorCode RDyCreatePrognosticVec(RDy rdy, Vec *prog_vec) {
  RDYB_CHECK(rdy && prog_vec && rdy->mesh, RDYB_ERR_ARG, "NULL argument");
#ifndef RDYB_HAVE_PETSC
  return rdyb_vec_create(3 * (int64_t)rdy->num_owned, 3, 0, prog_vec);
#else
  return rdyb_fail(RDYB_ERR_SUP, "with PETSc, create the Vec from RDycore's DM");
#endif
}

PetscErrorCode RDySetInitialConditions(RDy rdy, Vec ic) {
  RDYB_CALL(check_setup(rdy));
  RDYB_CHECK(ic, RDYB_ERR_ARG, "NULL Vec");
  int64_t n;
#ifndef RDYB_HAVE_PETSC
  RDYB_CALL(rdyb_vec_get_local_size(ic, &n));
  RDYB_CHECK(n == 3 * (int64_t)rdy->num_owned, RDYB_ERR_ARG_SIZ, "initial condition Vec has %lld entries (should be %d x 3)", (long long)n, rdy->num_owned);
  double *u;
  RDYB_CALL(rdyb_vec_get_array(ic, &u));
  int ierr = rdyb_operator_set_state(rdy->op, u);
  RDYB_CALL(rdyb_vec_restore_array(ic, &u));
  return ierr;
#else
  (void)n;
  return rdyb_fail(RDYB_ERR_SUP, "with PETSc, bind through VecGetArrayRead");
#endif
}
