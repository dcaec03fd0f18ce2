/* io_formats.c — Exodus II (NetCDF-3 classic / 64-bit offset) mesh reader and PETSc binary Vec reader/writer.
 * See include/rdyb200_io.h for what each entry point replaces in the reference.  Written against the published file formats
 * (NetCDF classic format specification; PETSc binary viewer: big-endian {VEC_FILE_CLASSID, n, n doubles}); the reference itself
 * reaches them through libexodus/libnetcdf and PETSc, none of which exist here. */
#define _XOPEN_SOURCE 700 /* pwrite, ftruncate */
#include "rdyb200_io.h"

#include <fcntl.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

#include "rdyb_error.h"

void rdyb_free(void *p) { free(p); }

static int slurp(const char *path, unsigned char **buf, size_t *len) {
  FILE *f = fopen(path, "rb");
  RDYB_CHECK(f, RDYB_ERR_USER, "Cannot open file %s", path);
  fseek(f, 0, SEEK_END);
  const long n = ftell(f);
  fseek(f, 0, SEEK_SET);
  if (n < 0) { fclose(f); return rdyb_fail(RDYB_ERR_USER, "Cannot size file %s", path); }
  *buf = (unsigned char *)malloc((size_t)n + 1);
  if (!*buf) { fclose(f); return rdyb_fail(RDYB_ERR_MEM, "out of memory reading %s", path); }
  const size_t got = fread(*buf, 1, (size_t)n, f);
  fclose(f);
  if (got != (size_t)n) { free(*buf); *buf = NULL; return rdyb_fail(RDYB_ERR_USER, "Short read of %s", path); }
  *len = (size_t)n;
  return RDYB_SUCCESS;
}

static uint32_t be32(const unsigned char *p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | (uint32_t)p[3]; }
static uint64_t be64(const unsigned char *p) { return ((uint64_t)be32(p) << 32) | be32(p + 4); }
static double bef64(const unsigned char *p) {
  const uint64_t u = be64(p);
  double d;
  memcpy(&d, &u, 8);
  return d;
}

/* ---------------- PETSc binary Vec ---------------- */
#define RDYB_VEC_FILE_CLASSID 1211214 /* 0x00127b4e */

int rdyb_petsc_vec_read(const char *path, int64_t *n, double **values, int32_t *index_bytes) {
  RDYB_CHECK(path && n && values, RDYB_ERR_ARG, "NULL argument");
  unsigned char *b = NULL;
  size_t len = 0;
  RDYB_CALL(slurp(path, &b, &len));
  size_t hdr = 0;
  int64_t count = -1;
  if (len >= 8 && be32(b) == RDYB_VEC_FILE_CLASSID) { hdr = 8; count = (int64_t)be32(b + 4); }
  else if (len >= 16 && be64(b) == RDYB_VEC_FILE_CLASSID) { hdr = 16; count = (int64_t)be64(b + 8); }
  if (count < 0 || (uint64_t)count > (len - hdr) / 8) { /* hdr <= len whenever count >= 0 */
    free(b);
    return rdyb_fail(RDYB_ERR_USER, "%s is not a PETSc binary Vec (no VEC_FILE_CLASSID header, or truncated)", path);
  }
  double *v = (double *)malloc(8 * (size_t)(count > 0 ? count : 1));
  if (!v) { free(b); return rdyb_fail(RDYB_ERR_MEM, "out of memory reading %s", path); }
  for (int64_t i = 0; i < count; ++i) v[i] = bef64(b + hdr + 8 * (size_t)i);
  free(b);
  *n = count; *values = v;
  if (index_bytes) *index_bytes = hdr == 8 ? 4 : 8;
  return RDYB_SUCCESS;
}

int rdyb_petsc_vec_write(const char *path, int64_t n, const double *values, int32_t index_bytes) {
  RDYB_CHECK(path && (values || n == 0) && n >= 0, RDYB_ERR_ARG, "bad argument");
  RDYB_CHECK(index_bytes == 4 || index_bytes == 8, RDYB_ERR_ARG, "index_bytes must be 4 or 8");
  FILE *f = fopen(path, "wb");
  RDYB_CHECK(f, RDYB_ERR_USER, "Cannot open file %s for writing", path);
  unsigned char h[16];
  const uint64_t ids[2] = {RDYB_VEC_FILE_CLASSID, (uint64_t)n};
  for (int k = 0; k < 2; ++k)
    for (int j = 0; j < index_bytes; ++j) h[k * index_bytes + j] = (unsigned char)(ids[k] >> (8 * (index_bytes - 1 - j)));
  int ok = fwrite(h, 1, 2 * (size_t)index_bytes, f) == 2 * (size_t)index_bytes;
  for (int64_t i = 0; ok && i < n; ++i) {
    uint64_t u;
    memcpy(&u, &values[i], 8);
    unsigned char d[8];
    for (int j = 0; j < 8; ++j) d[j] = (unsigned char)(u >> (8 * (7 - j)));
    ok = fwrite(d, 1, 8, f) == 8;
  }
  fclose(f);
  RDYB_CHECK(ok, RDYB_ERR_USER, "Short write to %s", path);
  return RDYB_SUCCESS;
}

/* ---------------- NetCDF-3 header ---------------- */
enum { NC_BYTE = 1, NC_CHAR = 2, NC_SHORT = 3, NC_INT = 4, NC_FLOAT = 5, NC_DOUBLE = 6 };
enum { NC_DIMENSION = 10, NC_VARIABLE = 11, NC_ATTRIBUTE = 12 };
#define NC_MAX_NAME 96
#define NC_MAX_DIMS 8

typedef struct { char name[NC_MAX_NAME]; int64_t len; } NcDim;
typedef struct {
  char name[NC_MAX_NAME];
  int ndims, dimids[NC_MAX_DIMS], type;
  int is_record;         /* first dimension is the unlimited one: data interleaved per record, not read here */
  int64_t nelems;        /* product of the dimension lengths */
  uint64_t begin;
  char elem_type[NC_MAX_NAME]; /* Exodus attribute of connect<k> */
} NcVar;
typedef struct {
  const unsigned char *b;
  size_t len, pos;
  int version, bad;
  int ndims, nvars;
  NcDim *dims;
  NcVar *vars;
} NcFile;

static uint32_t nc_u32(NcFile *f) {
  if (f->pos + 4 > f->len) { f->bad = 1; return 0; }
  const uint32_t v = be32(f->b + f->pos);
  f->pos += 4;
  return v;
}
static void nc_name(NcFile *f, char *out) {
  const uint32_t n = nc_u32(f);
  const size_t padded = ((size_t)n + 3) & ~(size_t)3;
  if (f->bad || f->pos + padded > f->len) { f->bad = 1; out[0] = 0; return; }
  const size_t keep = n < NC_MAX_NAME - 1 ? n : NC_MAX_NAME - 1;
  memcpy(out, f->b + f->pos, keep);
  out[keep] = 0;
  f->pos += padded;
}
static size_t nc_type_size(int t) { return t == NC_BYTE || t == NC_CHAR ? 1 : t == NC_SHORT ? 2 : t == NC_INT || t == NC_FLOAT ? 4 : t == NC_DOUBLE ? 8 : 0; }

/* att_list; copies the value of the CHAR attribute `want` (if given) into out */
static void nc_att_list(NcFile *f, const char *want, char *out) {
  const uint32_t tag = nc_u32(f), n = nc_u32(f);
  if (tag == 0 && n == 0) return;
  if (tag != NC_ATTRIBUTE) { f->bad = 1; return; }
  for (uint32_t i = 0; i < n && !f->bad; ++i) {
    char name[NC_MAX_NAME];
    nc_name(f, name);
    const int type = (int)nc_u32(f);
    const uint32_t cnt = nc_u32(f);
    const size_t bytes = (nc_type_size(type) * (size_t)cnt + 3) & ~(size_t)3;
    if (!nc_type_size(type) || f->pos + bytes > f->len) { f->bad = 1; return; }
    if (want && out && type == NC_CHAR && strcmp(name, want) == 0) {
      const size_t keep = cnt < NC_MAX_NAME - 1 ? cnt : NC_MAX_NAME - 1;
      memcpy(out, f->b + f->pos, keep);
      out[keep] = 0;
    }
    f->pos += bytes;
  }
}

static int nc_open(NcFile *f, const unsigned char *b, size_t len, const char *path) {
  memset(f, 0, sizeof(*f));
  f->b = b; f->len = len;
  RDYB_CHECK(len >= 8 && b[0] == 'C' && b[1] == 'D' && b[2] == 'F' && (b[3] == 1 || b[3] == 2), RDYB_ERR_USER,
             "%s is not a NetCDF-3 classic or 64-bit-offset file (Exodus II stored as HDF5/NetCDF-4 is not supported)", path);
  f->version = b[3];
  f->pos = 4;
  (void)nc_u32(f); /* numrecs */
  uint32_t tag = nc_u32(f), n = nc_u32(f);
  if (!(tag == 0 && n == 0)) {
    RDYB_CHECK(tag == NC_DIMENSION && n < 100000, RDYB_ERR_USER, "%s: corrupt NetCDF dimension list", path);
    f->dims = (NcDim *)calloc(n ? n : 1, sizeof(NcDim));
    RDYB_CHECK(f->dims, RDYB_ERR_MEM, "out of memory");
    f->ndims = (int)n;
    for (uint32_t i = 0; i < n; ++i) { nc_name(f, f->dims[i].name); f->dims[i].len = (int64_t)nc_u32(f); }
  }
  nc_att_list(f, NULL, NULL); /* global attributes */
  tag = nc_u32(f); n = nc_u32(f);
  if (!(tag == 0 && n == 0)) {
    RDYB_CHECK(!f->bad && tag == NC_VARIABLE && n < 100000, RDYB_ERR_USER, "%s: corrupt NetCDF variable list", path);
    f->vars = (NcVar *)calloc(n ? n : 1, sizeof(NcVar));
    RDYB_CHECK(f->vars, RDYB_ERR_MEM, "out of memory");
    f->nvars = (int)n;
    for (uint32_t i = 0; i < n && !f->bad; ++i) {
      NcVar *v = &f->vars[i];
      nc_name(f, v->name);
      const uint32_t nd = nc_u32(f);
      if (nd > NC_MAX_DIMS) { f->bad = 1; break; }
      v->ndims = (int)nd;
      v->nelems = 1;
      for (uint32_t d = 0; d < nd; ++d) {
        const uint32_t id = nc_u32(f);
        if (id >= (uint32_t)f->ndims) { f->bad = 1; break; }
        v->dimids[d] = (int)id;
        if (d == 0 && f->dims[id].len == 0) v->is_record = 1;
        else v->nelems *= f->dims[id].len;
      }
      nc_att_list(f, "elem_type", v->elem_type);
      v->type = (int)nc_u32(f);
      (void)nc_u32(f); /* vsize */
      if (f->version == 1) v->begin = nc_u32(f);
      else { const uint64_t hi = nc_u32(f); v->begin = (hi << 32) | nc_u32(f); }
    }
  }
  RDYB_CHECK(!f->bad, RDYB_ERR_USER, "%s: truncated or corrupt NetCDF header", path);
  return RDYB_SUCCESS;
}
static void nc_close(NcFile *f) { free(f->dims); free(f->vars); f->dims = NULL; f->vars = NULL; }
static int64_t nc_dim(const NcFile *f, const char *name) {
  for (int i = 0; i < f->ndims; ++i)
    if (strcmp(f->dims[i].name, name) == 0) return f->dims[i].len;
  return -1;
}
static const NcVar *nc_var(const NcFile *f, const char *name) {
  for (int i = 0; i < f->nvars; ++i)
    if (strcmp(f->vars[i].name, name) == 0) return &f->vars[i];
  return NULL;
}
/* a whole non-record variable, converted */
static int nc_read_double(const NcFile *f, const NcVar *v, double *out, int64_t expect, const char *path) {
  RDYB_CHECK(v && !v->is_record && v->nelems == expect && (v->type == NC_DOUBLE || v->type == NC_FLOAT), RDYB_ERR_USER, "%s: variable %s has an unexpected shape or type",
             path, v ? v->name : "(missing)");
  const size_t sz = nc_type_size(v->type);
  RDYB_CHECK(expect >= 0 && v->begin <= f->len && (uint64_t)expect <= (f->len - v->begin) / sz, RDYB_ERR_USER, "%s: variable %s runs past the end of the file", path,
             v->name);
  for (int64_t i = 0; i < expect; ++i) {
    const unsigned char *p = f->b + v->begin + sz * (size_t)i;
    if (v->type == NC_DOUBLE) out[i] = bef64(p);
    else { const uint32_t u = be32(p); float x; memcpy(&x, &u, 4); out[i] = x; }
  }
  return RDYB_SUCCESS;
}
static int nc_read_int(const NcFile *f, const NcVar *v, int32_t *out, int64_t expect, const char *path) {
  RDYB_CHECK(v && !v->is_record && v->nelems == expect && v->type == NC_INT, RDYB_ERR_USER, "%s: variable %s has an unexpected shape or type", path,
             v ? v->name : "(missing)");
  RDYB_CHECK(expect >= 0 && v->begin <= f->len && (uint64_t)expect <= (f->len - v->begin) / 4, RDYB_ERR_USER, "%s: variable %s runs past the end of the file", path,
             v->name);
  for (int64_t i = 0; i < expect; ++i) out[i] = (int32_t)be32(f->b + v->begin + 4 * (size_t)i);
  return RDYB_SUCCESS;
}

/* ---------------- Exodus II on top of NetCDF-3 ---------------- */
void rdyb_exodus_free(RdybExodusMesh *m) {
  if (!m) return;
  free(m->x); free(m->y); free(m->z); free(m->elem_offsets); free(m->elem_nodes); free(m->elem_block); free(m->block_ids);
  free(m->side_set_ids); free(m->side_set_offsets); free(m->side_elems); free(m->side_edges);
  memset(m, 0, sizeof(*m));
}

/* Exodus side number (1-based) of an element -> local edge (0-based), -1 for the two faces of a shell element.
 * 2-D QUAD/TRI: sides 1..n are the edges.  SHELL4, and TRI3/TRI in a 3-D file (triangular shells): sides 1, 2 are the faces and
 * 3..n+2 the edges — the numbering the side sets of share/meshes/{quad_tri_mesh,mms_triangles_dx1,Houston1km_with_z}.exo use. */
static int exodus_side_to_edge(const char *elem_type, int nodes_per_elem, int num_dim, int side) {
  const int shell = strncmp(elem_type, "SHELL", 5) == 0 || strncmp(elem_type, "shell", 5) == 0 ||
                    ((strncmp(elem_type, "TRI", 3) == 0 || strncmp(elem_type, "tri", 3) == 0) && num_dim == 3);
  const int64_t e = shell ? (int64_t)side - 3 : (int64_t)side - 1;
  return (e >= 0 && e < nodes_per_elem) ? (int)e : -1;
}

static int exodus_build(const NcFile *f, RdybExodusMesh *m, const char *path) {
  const int64_t nn = nc_dim(f, "num_nodes"), ne = nc_dim(f, "num_elem"), nb = nc_dim(f, "num_el_blk");
  int64_t nd = nc_dim(f, "num_dim"), nss = nc_dim(f, "num_side_sets");
  RDYB_CHECK(nn > 0 && ne > 0 && nb > 0 && nd >= 2 && nd <= 3, RDYB_ERR_USER, "%s is not an Exodus II mesh (num_nodes/num_elem/num_el_blk/num_dim missing)", path);
  if (nss < 0) nss = 0;
  /* sizes a file of this length can hold (>= 4 bytes per coordinate and per connectivity entry; one dimension per block and per
   * side set): a corrupt header must not drive the allocations and loops below */
  RDYB_CHECK((uint64_t)nn <= f->len / 4 && (uint64_t)ne <= f->len / 12 && nb <= f->ndims && nss <= f->ndims, RDYB_ERR_USER,
             "%s: corrupt Exodus II header (%lld nodes, %lld elements, %lld blocks, %lld side sets in a file of %zu bytes)", path, (long long)nn, (long long)ne,
             (long long)nb, (long long)nss, f->len);
  m->num_dim = (int32_t)nd; m->num_nodes = (int32_t)nn; m->num_elems = (int32_t)ne; m->num_blocks = (int32_t)nb; m->num_side_sets = (int32_t)nss;
  m->x = (double *)calloc((size_t)nn, 8); m->y = (double *)calloc((size_t)nn, 8); m->z = (double *)calloc((size_t)nn, 8);
  m->elem_offsets = (int32_t *)calloc((size_t)ne + 1, 4); m->elem_block = (int32_t *)calloc((size_t)ne, 4); m->block_ids = (int32_t *)calloc((size_t)nb, 4);
  m->side_set_ids = (int32_t *)calloc((size_t)nss + 1, 4); m->side_set_offsets = (int32_t *)calloc((size_t)nss + 1, 4);
  RDYB_CHECK(m->x && m->y && m->z && m->elem_offsets && m->elem_block && m->block_ids && m->side_set_ids && m->side_set_offsets, RDYB_ERR_MEM, "out of memory");
  /* coordinates: coordx/coordy/coordz, or the older single "coord"[num_dim][num_nodes] */
  if (nc_var(f, "coordx")) {
    RDYB_CALL(nc_read_double(f, nc_var(f, "coordx"), m->x, nn, path));
    RDYB_CALL(nc_read_double(f, nc_var(f, "coordy"), m->y, nn, path));
    if (nd == 3) RDYB_CALL(nc_read_double(f, nc_var(f, "coordz"), m->z, nn, path));
  } else {
    double *c = (double *)malloc(8 * (size_t)(nd * nn));
    RDYB_CHECK(c, RDYB_ERR_MEM, "out of memory");
    const int rc = nc_read_double(f, nc_var(f, "coord"), c, nd * nn, path);
    if (rc) { free(c); return rc; }
    memcpy(m->x, c, 8 * (size_t)nn); memcpy(m->y, c + nn, 8 * (size_t)nn);
    if (nd == 3) memcpy(m->z, c + 2 * nn, 8 * (size_t)nn);
    free(c);
  }
  if (nc_var(f, "eb_prop1")) RDYB_CALL(nc_read_int(f, nc_var(f, "eb_prop1"), m->block_ids, nb, path));
  else for (int64_t b = 0; b < nb; ++b) m->block_ids[b] = (int32_t)b + 1;
  /* element blocks, in file order */
  int64_t total = 0;
  char name[64];
  for (int64_t b = 0; b < nb; ++b) {
    snprintf(name, sizeof(name), "num_el_in_blk%lld", (long long)b + 1);
    const int64_t nel = nc_dim(f, name);
    snprintf(name, sizeof(name), "num_nod_per_el%lld", (long long)b + 1);
    const int64_t npe = nc_dim(f, name);
    RDYB_CHECK(nel >= 0 && npe >= 3 && npe <= 4, RDYB_ERR_USER, "%s: element block %lld has %lld nodes per element (only triangles and quadrilaterals are supported)",
               path, (long long)b + 1, (long long)npe);
    RDYB_CHECK(nel <= ne, RDYB_ERR_USER, "%s: element block %lld holds %lld elements, num_elem says %lld", path, (long long)b + 1, (long long)nel, (long long)ne);
    total += nel * npe;
  }
  RDYB_CHECK(total <= 4 * ne, RDYB_ERR_USER, "%s: element blocks hold more elements than num_elem (%lld)", path, (long long)ne);
  m->elem_nodes = (int32_t *)calloc((size_t)(total > 0 ? total : 1), 4);
  char *et8 = (char *)calloc((size_t)ne, 8); /* first 7 characters of the element type, per element */
  RDYB_CHECK(m->elem_nodes && et8, RDYB_ERR_MEM, "out of memory");
  int64_t e0 = 0, pos = 0;
  int rc = RDYB_SUCCESS;
  for (int64_t b = 0; b < nb && !rc; ++b) {
    snprintf(name, sizeof(name), "num_el_in_blk%lld", (long long)b + 1);
    const int64_t nel = nc_dim(f, name);
    snprintf(name, sizeof(name), "num_nod_per_el%lld", (long long)b + 1);
    const int64_t npe = nc_dim(f, name);
    snprintf(name, sizeof(name), "connect%lld", (long long)b + 1);
    const NcVar *cv = nc_var(f, name);
    if (nel == 0) continue;
    if (e0 + nel > ne) { rc = rdyb_fail(RDYB_ERR_USER, "%s: element blocks hold more than the %lld elements num_elem says", path, (long long)ne); break; }
    rc = nc_read_int(f, cv, m->elem_nodes + pos, nel * npe, path);
    if (rc) break;
    for (int64_t k = 0; k < nel * npe; ++k) {
      const int64_t v = (int64_t)m->elem_nodes[pos + k] - 1; /* Exodus is 1-based */
      if (v < 0 || v >= nn) { rc = rdyb_fail(RDYB_ERR_USER, "%s: %s refers to node %lld of %lld", path, name, (long long)v + 1, (long long)nn); break; }
      m->elem_nodes[pos + k] = (int32_t)v;
    }
    for (int64_t k = 0; k < nel; ++k) {
      m->elem_offsets[e0 + k] = (int32_t)(pos + k * npe);
      m->elem_block[e0 + k] = m->block_ids[b];
      memcpy(et8 + 8 * (e0 + k), cv->elem_type, 7);
    }
    e0 += nel; pos += nel * npe;
  }
  if (!rc && e0 != ne) rc = rdyb_fail(RDYB_ERR_USER, "%s: element blocks hold %lld elements, num_elem says %lld", path, (long long)e0, (long long)ne);
  m->elem_offsets[ne] = (int32_t)pos;
  /* side sets */
  if (!rc && nss > 0) {
    if (nc_var(f, "ss_prop1")) rc = nc_read_int(f, nc_var(f, "ss_prop1"), m->side_set_ids, nss, path);
    else for (int64_t s = 0; s < nss; ++s) m->side_set_ids[s] = (int32_t)s + 1;
    int64_t nsides = 0;
    for (int64_t s = 0; s < nss && !rc; ++s) {
      snprintf(name, sizeof(name), "num_side_ss%lld", (long long)s + 1);
      const int64_t k = nc_dim(f, name);
      m->side_set_offsets[s] = (int32_t)nsides;
      nsides += k > 0 ? k : 0;
      if ((uint64_t)nsides > f->len / 8) rc = rdyb_fail(RDYB_ERR_USER, "%s: corrupt side set sizes (%lld sides in a file of %zu bytes)", path, (long long)nsides, f->len);
    }
    if (rc) nsides = 0;
    m->side_set_offsets[nss] = (int32_t)nsides;
    m->side_elems = (int32_t *)calloc((size_t)(nsides > 0 ? nsides : 1), 4);
    m->side_edges = (int32_t *)calloc((size_t)(nsides > 0 ? nsides : 1), 4);
    if (!rc && (!m->side_elems || !m->side_edges)) rc = rdyb_fail(RDYB_ERR_MEM, "out of memory");
    for (int64_t s = 0; s < nss && !rc; ++s) {
      const int64_t k = m->side_set_offsets[s + 1] - m->side_set_offsets[s], o = m->side_set_offsets[s];
      if (k == 0) continue;
      snprintf(name, sizeof(name), "elem_ss%lld", (long long)s + 1);
      rc = nc_read_int(f, nc_var(f, name), m->side_elems + o, k, path);
      if (rc) break;
      snprintf(name, sizeof(name), "side_ss%lld", (long long)s + 1);
      rc = nc_read_int(f, nc_var(f, name), m->side_edges + o, k, path);
      for (int64_t i = 0; i < k && !rc; ++i) {
        const int64_t el = (int64_t)m->side_elems[o + i] - 1;
        if (el < 0 || el >= ne) { rc = rdyb_fail(RDYB_ERR_USER, "%s: side set %lld refers to element %lld of %lld", path, (long long)s + 1, (long long)el + 1, (long long)ne); break; }
        m->side_elems[o + i] = (int32_t)el;
        m->side_edges[o + i] = exodus_side_to_edge(et8 + 8 * el, m->elem_offsets[el + 1] - m->elem_offsets[el], (int)nd, m->side_edges[o + i]);
      }
    }
  }
  free(et8);
  return rc;
}

int rdyb_exodus_read(const char *path, RdybExodusMesh *mesh) {
  RDYB_CHECK(path && mesh, RDYB_ERR_ARG, "NULL argument");
  memset(mesh, 0, sizeof(*mesh));
  unsigned char *b = NULL;
  size_t len = 0;
  RDYB_CALL(slurp(path, &b, &len));
  NcFile f;
  int rc = nc_open(&f, b, len, path);
  if (!rc) rc = exodus_build(&f, mesh, path);
  nc_close(&f);
  free(b);
  if (rc) rdyb_exodus_free(mesh);
  return rc;
}

/* ---------------- checkpoints: PetscBag "metadata" + natural-order Vec (src/checkpoint.c) ---------------- */
#define RDYB_BAG_FILE_CLASSID 1211219
#define RDYB_BAG_NAME_LENGTH 64
#define RDYB_BAG_HELP_LENGTH 128
enum { RDYB_PETSC_DOUBLE = 1, RDYB_PETSC_INT = 16 }; /* PetscDataType */
/* what an item offset would be in a 64-bit PETSc build: the bag header rounded up to a PetscScalar, then the field's offset in
 * CheckpointMetadata (src/checkpoint.c:49-54).  Informational: PetscBagLoad walks the registered items, not these numbers. */
#define RDYB_BAG_HEADER_BYTES 248

int rdyb_checkpoint_filename_base(const char *directory, const char *prefix, int64_t index, int64_t max_index, char *out, int32_t capacity) {
  RDYB_CHECK(directory && prefix && out && capacity > 0, RDYB_ERR_ARG, "NULL argument");
  RDYB_CHECK(index >= 0 && max_index >= 1, RDYB_ERR_ARG, "checkpoint index %lld of %lld", (long long)index, (long long)max_index);
  int digits = 1;
  for (int64_t v = max_index; v >= 10; v /= 10) ++digits;
  const int n = snprintf(out, (size_t)capacity, "%s/%s.rdycore.r.%0*lld", directory, prefix, digits, (long long)index);
  RDYB_CHECK(n > 0 && n < capacity, RDYB_ERR_ARG_SIZ, "checkpoint file name does not fit %d bytes", capacity);
  return RDYB_SUCCESS;
}

static size_t put_int(unsigned char *p, int64_t v, int index_bytes) {
  for (int j = 0; j < index_bytes; ++j) p[j] = (unsigned char)((uint64_t)v >> (8 * (index_bytes - 1 - j)));
  return (size_t)index_bytes;
}
static size_t put_f64(unsigned char *p, double d) {
  uint64_t u;
  memcpy(&u, &d, 8);
  for (int j = 0; j < 8; ++j) p[j] = (unsigned char)(u >> (8 * (7 - j)));
  return 8;
}
static size_t put_str(unsigned char *p, const char *s, size_t field) {
  memset(p, 0, field);
  strncpy((char *)p, s, field - 1);
  return field;
}
static size_t bag_item_header(unsigned char *p, int64_t offset, int dtype, const char *name, const char *help, int index_bytes) {
  size_t k = 0;
  k += put_int(p + k, RDYB_BAG_HEADER_BYTES + offset, index_bytes);
  k += put_int(p + k, dtype, index_bytes);
  k += put_str(p + k, name, RDYB_BAG_NAME_LENGTH);
  k += put_str(p + k, help, RDYB_BAG_HELP_LENGTH);
  k += put_int(p + k, 1, index_bytes); /* msize: one entry */
  return k;
}
/* bytes of the bag record: 3 ints + name + help, then 4 items of {3 ints + name + help + value} */
static size_t bag_bytes(int index_bytes) {
  const size_t item = 3 * (size_t)index_bytes + RDYB_BAG_NAME_LENGTH + RDYB_BAG_HELP_LENGTH;
  return item + 4 * item + 2 * (size_t)index_bytes + 2 * 8;
}

int rdyb_checkpoint_write(const char *path, const RdybCheckpointMetadata *md, int64_t num_global_cells, int32_t ndof, int32_t index_bytes,
                          int64_t n_local, const int64_t *natural_ids, const double *values, int32_t write_header) {
  RDYB_CHECK(path && (values || n_local == 0), RDYB_ERR_ARG, "NULL argument");
  RDYB_CHECK(index_bytes == 4 || index_bytes == 8, RDYB_ERR_ARG, "index_bytes must be 4 or 8");
  RDYB_CHECK(num_global_cells >= 0 && ndof >= 1 && n_local >= 0 && n_local <= num_global_cells, RDYB_ERR_ARG, "bad sizes");
  RDYB_CHECK(!write_header || md, RDYB_ERR_ARG, "the header needs the metadata");
  const size_t bag = bag_bytes(index_bytes), data0 = bag + 2 * (size_t)index_bytes;
  const int fd = open(path, O_WRONLY | O_CREAT, 0644);
  RDYB_CHECK(fd >= 0, RDYB_ERR_USER, "Cannot open file %s for writing", path);
  int ok = 1;
  if (write_header) {
    unsigned char *h = (unsigned char *)malloc(data0);
    if (!h) { close(fd); return rdyb_fail(RDYB_ERR_MEM, "out of memory"); }
    size_t k = 0;
    k += put_int(h + k, RDYB_BAG_FILE_CLASSID, index_bytes);
    k += put_int(h + k, 0, index_bytes); /* deprecated bag size */
    k += put_int(h + k, 4, index_bytes); /* items */
    k += put_str(h + k, "metadata", RDYB_BAG_NAME_LENGTH);
    k += put_str(h + k, "Checkpoint metadata", RDYB_BAG_HELP_LENGTH);
    /* field offsets in {PetscInt nproc; PetscReal t; PetscReal dt; PetscInt step;}: 0, 8, 16, 24 for either PetscInt size */
    k += bag_item_header(h + k, 0, RDYB_PETSC_INT, "nproc", "Number of MPI tasks", index_bytes);
    k += put_int(h + k, md->nproc, index_bytes);
    k += bag_item_header(h + k, 8, RDYB_PETSC_DOUBLE, "t", "Simulation time", index_bytes);
    k += put_f64(h + k, md->t);
    k += bag_item_header(h + k, 16, RDYB_PETSC_DOUBLE, "dt", "Simulation timestep", index_bytes);
    k += put_f64(h + k, md->dt);
    k += bag_item_header(h + k, 24, RDYB_PETSC_INT, "step", "Simulation step number", index_bytes);
    k += put_int(h + k, md->step, index_bytes);
    k += put_int(h + k, RDYB_VEC_FILE_CLASSID, index_bytes);
    k += put_int(h + k, num_global_cells * ndof, index_bytes);
    ok = k == data0 && pwrite(fd, h, data0, 0) == (ssize_t)data0 && ftruncate(fd, (off_t)(data0 + 8 * (size_t)(num_global_cells * ndof))) == 0;
    free(h);
    if (ok) { /* the viewer's .info file: the block size VecLoad picks up */
      char info[4096];
      if (snprintf(info, sizeof info, "%s.info", path) < (int)sizeof info) {
        FILE *f = fopen(info, "w");
        if (f) { fprintf(f, "-vecload_block_size %d\n", ndof); fclose(f); }
      }
    }
  }
  /* rows: runs of consecutive natural ids go out in one pwrite */
  const size_t row = 8 * (size_t)ndof, cap = 4096;
  unsigned char *buf = (unsigned char *)malloc(cap * row);
  if (!buf) { close(fd); return rdyb_fail(RDYB_ERR_MEM, "out of memory"); }
  for (int64_t i = 0; ok && i < n_local;) {
    const int64_t id0 = natural_ids ? natural_ids[i] : i;
    if (id0 < 0 || id0 >= num_global_cells) { ok = 0; rdyb_fail(RDYB_ERR_USER, "natural cell id %lld out of range", (long long)id0); free(buf); close(fd); return RDYB_ERR_USER; }
    int64_t j = i + 1;
    while (j < n_local && (size_t)(j - i) < cap && (natural_ids ? natural_ids[j] : j) == id0 + (j - i)) ++j;
    for (int64_t r = i; r < j; ++r)
      for (int c = 0; c < ndof; ++c) put_f64(buf + (size_t)(r - i) * row + 8 * (size_t)c, values[(size_t)r * ndof + c]);
    const size_t nb = (size_t)(j - i) * row;
    ok = pwrite(fd, buf, nb, (off_t)(data0 + (size_t)id0 * row)) == (ssize_t)nb;
    i = j;
  }
  free(buf);
  ok = close(fd) == 0 && ok;
  RDYB_CHECK(ok, RDYB_ERR_USER, "Short write to %s", path);
  return RDYB_SUCCESS;
}

int rdyb_checkpoint_read(const char *path, RdybCheckpointMetadata *md, int64_t *n_values, double **values, int32_t *index_bytes) {
  RDYB_CHECK(path && md && n_values && values, RDYB_ERR_ARG, "NULL argument");
  RDYB_CHECK(strstr(path, ".bin"), RDYB_ERR_USER, "Invalid checkpoint file: %s", path); /* src/checkpoint.c:271-280; .h5 is not read here */
  unsigned char *b = NULL;
  size_t len = 0;
  RDYB_CALL(slurp(path, &b, &len));
  int ib = 0;
  if (len >= 8 && be32(b) == RDYB_BAG_FILE_CLASSID) ib = 4;
  else if (len >= 16 && be64(b) == RDYB_BAG_FILE_CLASSID) ib = 8;
#define GET_INT(at) (ib == 4 ? (int64_t)(int32_t)be32(b + (at)) : (int64_t)be64(b + (at)))
  int ierr = RDYB_SUCCESS;
  if (!ib || len < bag_bytes(ib) + 2 * (size_t)ib || GET_INT(2 * (size_t)ib) != 4) ierr = rdyb_fail(RDYB_ERR_USER, "%s does not start with the checkpoint metadata bag", path);
  if (!ierr) {
    const size_t item = 3 * (size_t)ib + RDYB_BAG_NAME_LENGTH + RDYB_BAG_HELP_LENGTH;
    static const char *names[4] = {"nproc", "t", "dt", "step"};
    size_t k = item;
    for (int i = 0; i < 4 && !ierr; ++i) { /* the items in the order CreateMetadata registers them */
      const int64_t dtype = GET_INT(k + (size_t)ib), msize = GET_INT(k + 2 * (size_t)ib + RDYB_BAG_NAME_LENGTH + RDYB_BAG_HELP_LENGTH);
      const int is_int = i == 0 || i == 3;
      if (strncmp((const char *)b + k + 2 * (size_t)ib, names[i], RDYB_BAG_NAME_LENGTH) != 0 || msize != 1 || dtype != (is_int ? RDYB_PETSC_INT : RDYB_PETSC_DOUBLE)) {
        ierr = rdyb_fail(RDYB_ERR_USER, "%s: metadata item %d is not '%s'", path, i, names[i]);
        break;
      }
      k += item;
      if (i == 0) md->nproc = (int32_t)GET_INT(k);
      else if (i == 1) md->t = bef64(b + k);
      else if (i == 2) md->dt = bef64(b + k);
      else md->step = (int32_t)GET_INT(k);
      k += is_int ? (size_t)ib : 8;
    }
    if (!ierr) {
      const int64_t count = GET_INT(k) == RDYB_VEC_FILE_CLASSID ? GET_INT(k + (size_t)ib) : -1;
      k += 2 * (size_t)ib;
      if (count < 0 || k > len || (uint64_t)count > (len - k) / 8) ierr = rdyb_fail(RDYB_ERR_USER, "%s: no solution Vec after the metadata, or truncated", path);
      else {
        double *v = (double *)malloc(8 * (size_t)(count > 0 ? count : 1));
        if (!v) ierr = rdyb_fail(RDYB_ERR_MEM, "out of memory reading %s", path);
        else {
          for (int64_t i = 0; i < count; ++i) v[i] = bef64(b + k + 8 * (size_t)i);
          *n_values = count; *values = v;
          if (index_bytes) *index_bytes = ib;
        }
      }
    }
  }
#undef GET_INT
  free(b);
  return ierr;
}
