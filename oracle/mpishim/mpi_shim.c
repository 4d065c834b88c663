/* TEST INFRASTRUCTURE — single-rank implementation of oracle/mpishim/mpi.h (see there). */
#define _GNU_SOURCE
#include "mpi.h"
#include <fcntl.h>
#include <sys/stat.h>
#include <time.h>
#include <unistd.h>

#define B(sz) {sz, sz, 0, 1, {0}, {0}}
struct kmi_mpi_type kmi_mpi_builtin[32] = {
  B(1), B(1), B(1), B(1), B(2), B(2), B(4), B(4), B(8), B(8), B(8), B(8), B(4), B(8), B(16),
  B(8), B(16), B(16), B(8), B(8), B(32), B(4), B(1)};
static int g_init = 0, g_fin = 0, g_keyvals = 0;

int MPI_Init(int* a, char*** b) { (void)a; (void)b; g_init = 1; return MPI_SUCCESS; }
int MPI_Init_thread(int* a, char*** b, int req, int* prov) { (void)a; (void)b; if (prov) *prov = req; g_init = 1; return MPI_SUCCESS; }
int MPI_Finalize(void) { g_fin = 1; return MPI_SUCCESS; }
int MPI_Initialized(int* f) { *f = g_init; return MPI_SUCCESS; }
int MPI_Finalized(int* f) { *f = g_fin; return MPI_SUCCESS; }
int MPI_Abort(MPI_Comm c, int code) { (void)c; fprintf(stderr, "MPI_Abort(%d)\n", code); exit(code ? code : 1); }
int MPI_Comm_rank(MPI_Comm c, int* r) { (void)c; *r = 0; return MPI_SUCCESS; }
int MPI_Comm_size(MPI_Comm c, int* s) { (void)c; *s = 1; return MPI_SUCCESS; }
int MPI_Comm_dup(MPI_Comm c, MPI_Comm* n) { *n = c ? c + 16 : 17; return MPI_SUCCESS; }
int MPI_Comm_free(MPI_Comm* c) { *c = MPI_COMM_NULL; return MPI_SUCCESS; }
int MPI_Comm_split(MPI_Comm c, int color, int key, MPI_Comm* n) { (void)key; *n = color == MPI_UNDEFINED ? MPI_COMM_NULL : c + 16; return MPI_SUCCESS; }
int MPI_Comm_split_type(MPI_Comm c, int t, int key, MPI_Info i, MPI_Comm* n) { (void)t; (void)key; (void)i; *n = c + 16; return MPI_SUCCESS; }
int MPI_Comm_compare(MPI_Comm a, MPI_Comm b, int* r) { *r = a == b ? 0 : 1; return MPI_SUCCESS; }
int MPI_Comm_set_errhandler(MPI_Comm c, MPI_Errhandler e) { (void)c; (void)e; return MPI_SUCCESS; }
int MPI_Comm_create_errhandler(MPI_Comm_errhandler_function* f, MPI_Errhandler* e) { (void)f; *e = 3; return MPI_SUCCESS; }
int MPI_Errhandler_create(MPI_Handler_function* f, MPI_Errhandler* e) { (void)f; *e = 3; return MPI_SUCCESS; }
int MPI_Errhandler_set(MPI_Comm c, MPI_Errhandler e) { (void)c; (void)e; return MPI_SUCCESS; }
int MPI_Errhandler_free(MPI_Errhandler* e) { *e = 0; return MPI_SUCCESS; }
int MPI_Error_string(int code, char* s, int* len) { *len = snprintf(s, MPI_MAX_ERROR_STRING, "mpi shim error %d", code); return MPI_SUCCESS; }
int MPI_Error_class(int code, int* cls) { *cls = code; return MPI_SUCCESS; }
int MPI_Get_processor_name(char* s, int* len) { *len = snprintf(s, MPI_MAX_PROCESSOR_NAME, "localhost"); return MPI_SUCCESS; }
int MPI_Barrier(MPI_Comm c) { (void)c; return MPI_SUCCESS; }
double MPI_Wtime(void) { struct timespec t; clock_gettime(CLOCK_MONOTONIC, &t); return t.tv_sec + 1e-9 * t.tv_nsec; }

static MPI_Datatype mk(long size, long extent, long lb) {
  MPI_Datatype t = (MPI_Datatype)calloc(1, sizeof(*t));
  t->size = size; t->extent = extent; t->lb = lb;
  return t;
}
int MPI_Type_contiguous(int n, MPI_Datatype o, MPI_Datatype* t) { *t = mk(n * o->size, n * o->extent, o->lb); return MPI_SUCCESS; }
int MPI_Type_vector(int n, int bl, int stride, MPI_Datatype o, MPI_Datatype* t) {
  *t = mk((long)n * bl * o->size, n ? ((long)(n - 1) * stride + bl) * o->extent : 0, o->lb); return MPI_SUCCESS; }
int MPI_Type_create_struct(int n, const int* bl, const MPI_Aint* disp, const MPI_Datatype* ts, MPI_Datatype* t) {
  long size = 0, lo = 0, hi = 0;
  for (int i = 0; i < n; ++i) {
    size += bl[i] * ts[i]->size;
    long b = disp[i] + ts[i]->lb, e = disp[i] + ts[i]->lb + bl[i] * ts[i]->extent;
    if (i == 0 || b < lo) lo = b;
    if (i == 0 || e > hi) hi = e;
  }
  *t = mk(size, hi - lo, lo); return MPI_SUCCESS; }
int MPI_Type_create_resized(MPI_Datatype o, MPI_Aint lb, MPI_Aint extent, MPI_Datatype* t) { *t = mk(o->size, extent, lb); return MPI_SUCCESS; }
int MPI_Type_dup(MPI_Datatype o, MPI_Datatype* t) { *t = mk(o->size, o->extent, o->lb); memcpy((*t)->attr, o->attr, sizeof(o->attr)); memcpy((*t)->attr_set, o->attr_set, sizeof(o->attr_set)); return MPI_SUCCESS; }
int MPI_Type_commit(MPI_Datatype* t) { (void)t; return MPI_SUCCESS; }
int MPI_Type_free(MPI_Datatype* t) { if (*t && !(*t)->builtin) free(*t); *t = MPI_DATATYPE_NULL; return MPI_SUCCESS; }
int MPI_Type_size(MPI_Datatype t, int* s) { *s = (int)t->size; return MPI_SUCCESS; }
int MPI_Type_get_extent(MPI_Datatype t, MPI_Aint* lb, MPI_Aint* e) { *lb = t->lb; *e = t->extent; return MPI_SUCCESS; }
int MPI_Type_get_true_extent(MPI_Datatype t, MPI_Aint* lb, MPI_Aint* e) { *lb = t->lb; *e = t->extent; return MPI_SUCCESS; }
int MPI_Type_create_keyval(MPI_Type_copy_attr_function* c, MPI_Type_delete_attr_function* d, int* k, void* x) { (void)c; (void)d; (void)x; *k = g_keyvals++ & 7; return MPI_SUCCESS; }
int MPI_Type_set_attr(MPI_Datatype t, int k, void* v) { t->attr[k & 7] = v; t->attr_set[k & 7] = 1; return MPI_SUCCESS; }
int MPI_Type_get_attr(MPI_Datatype t, int k, void* v, int* flag) { *flag = t->attr_set[k & 7]; if (*flag) *(void**)v = t->attr[k & 7]; return MPI_SUCCESS; }
int MPI_Get_address(const void* p, MPI_Aint* a) { *a = (MPI_Aint)p; return MPI_SUCCESS; }
int MPI_Get_count(const MPI_Status* s, MPI_Datatype t, int* n) { *n = t->size ? (int)(s->count_ / t->size) : 0; return MPI_SUCCESS; }
int MPI_Get_elements_x(const MPI_Status* s, MPI_Datatype t, MPI_Count* n) { *n = t->size ? s->count_ / t->size : 0; return MPI_SUCCESS; }
int MPI_Op_create(MPI_User_function* f, int commute, MPI_Op* op) { (void)f; (void)commute; *op = MPI_OP_USER_; return MPI_SUCCESS; }
int MPI_Op_free(MPI_Op* op) { *op = MPI_OP_NULL; return MPI_SUCCESS; }

/* n elements of type t laid out with the type's extent */
static void cp(const void* s, void* d, long n, MPI_Datatype t) { if (s != MPI_IN_PLACE && s != d && n > 0) memmove(d, s, (size_t)(n * t->extent)); }
int MPI_Bcast(void* b, int n, MPI_Datatype t, int root, MPI_Comm c) { (void)b; (void)n; (void)t; (void)root; (void)c; return MPI_SUCCESS; }
int MPI_Reduce(const void* s, void* r, int n, MPI_Datatype t, MPI_Op op, int root, MPI_Comm c) { (void)op; (void)root; (void)c; cp(s, r, n, t); return MPI_SUCCESS; }
int MPI_Allreduce(const void* s, void* r, int n, MPI_Datatype t, MPI_Op op, MPI_Comm c) { (void)op; (void)c; cp(s, r, n, t); return MPI_SUCCESS; }
int MPI_Scan(const void* s, void* r, int n, MPI_Datatype t, MPI_Op op, MPI_Comm c) { (void)op; (void)c; cp(s, r, n, t); return MPI_SUCCESS; }
int MPI_Exscan(const void* s, void* r, int n, MPI_Datatype t, MPI_Op op, MPI_Comm c) { (void)s; (void)r; (void)n; (void)t; (void)op; (void)c; return MPI_SUCCESS; }   /* rank 0: recvbuf undefined */
int MPI_Gather(const void* s, int sn, MPI_Datatype st, void* r, int rn, MPI_Datatype rt, int root, MPI_Comm c) { (void)rn; (void)rt; (void)root; (void)c; cp(s, r, sn, st); return MPI_SUCCESS; }
int MPI_Gatherv(const void* s, int sn, MPI_Datatype st, void* r, const int* rc, const int* d, MPI_Datatype rt, int root, MPI_Comm c) {
  (void)rc; (void)root; (void)c; if (s != MPI_IN_PLACE) memmove((char*)r + (d ? d[0] : 0) * rt->extent, s, (size_t)(sn * st->extent)); return MPI_SUCCESS; }
int MPI_Allgather(const void* s, int sn, MPI_Datatype st, void* r, int rn, MPI_Datatype rt, MPI_Comm c) { (void)rn; (void)rt; (void)c; cp(s, r, sn, st); return MPI_SUCCESS; }
int MPI_Allgatherv(const void* s, int sn, MPI_Datatype st, void* r, const int* rc, const int* d, MPI_Datatype rt, MPI_Comm c) { return MPI_Gatherv(s, sn, st, r, rc, d, rt, 0, c); }
int MPI_Scatter(const void* s, int sn, MPI_Datatype st, void* r, int rn, MPI_Datatype rt, int root, MPI_Comm c) { (void)rn; (void)rt; (void)root; (void)c; if (r != MPI_IN_PLACE) cp(s, r, sn, st); return MPI_SUCCESS; }
int MPI_Scatterv(const void* s, const int* sc, const int* d, MPI_Datatype st, void* r, int rn, MPI_Datatype rt, int root, MPI_Comm c) {
  (void)rn; (void)rt; (void)root; (void)c; if (r != MPI_IN_PLACE) memmove(r, (const char*)s + (d ? d[0] : 0) * st->extent, (size_t)(sc[0] * st->extent)); return MPI_SUCCESS; }
int MPI_Alltoall(const void* s, int sn, MPI_Datatype st, void* r, int rn, MPI_Datatype rt, MPI_Comm c) { (void)rn; (void)rt; (void)c; cp(s, r, sn, st); return MPI_SUCCESS; }
int MPI_Alltoallv(const void* s, const int* sc, const int* sd, MPI_Datatype st, void* r, const int* rc, const int* rd, MPI_Datatype rt, MPI_Comm c) {
  (void)rc; (void)c; if (s != MPI_IN_PLACE) memmove((char*)r + (rd ? rd[0] : 0) * rt->extent, (const char*)s + (sd ? sd[0] : 0) * st->extent, (size_t)(sc[0] * st->extent)); return MPI_SUCCESS; }
int MPI_Alltoallw(const void* s, const int* sc, const int* sd, const MPI_Datatype* st, void* r, const int* rc, const int* rd, const MPI_Datatype* rt, MPI_Comm c) {
  (void)rc; (void)rt; (void)c; if (s != MPI_IN_PLACE) memmove((char*)r + (rd ? rd[0] : 0), (const char*)s + (sd ? sd[0] : 0), (size_t)(sc[0] * st[0]->extent)); return MPI_SUCCESS; }

/* one-message mailbox to self */
static void* g_box = 0; static long g_box_n = 0; static int g_box_tag = 0;
int MPI_Send(const void* b, int n, MPI_Datatype t, int dst, int tag, MPI_Comm c) { (void)dst; (void)c; free(g_box); g_box_n = n * t->extent; g_box = malloc(g_box_n ? g_box_n : 1); memcpy(g_box, b, g_box_n); g_box_tag = tag; return MPI_SUCCESS; }
int MPI_Recv(void* b, int n, MPI_Datatype t, int src, int tag, MPI_Comm c, MPI_Status* s) {
  (void)src; (void)tag; (void)c; long m = n * t->extent < g_box_n ? n * t->extent : g_box_n; if (g_box) memcpy(b, g_box, m);
  if (s) { s->MPI_SOURCE = 0; s->MPI_TAG = g_box_tag; s->MPI_ERROR = 0; s->count_ = m; } return MPI_SUCCESS; }
int MPI_Isend(const void* b, int n, MPI_Datatype t, int dst, int tag, MPI_Comm c, MPI_Request* r) { *r = 1; return MPI_Send(b, n, t, dst, tag, c); }
static void* g_irecv_buf = 0; static long g_irecv_n = 0;
int MPI_Irecv(void* b, int n, MPI_Datatype t, int src, int tag, MPI_Comm c, MPI_Request* r) { (void)src; (void)tag; (void)c; g_irecv_buf = b; g_irecv_n = n * t->extent; *r = 2; return MPI_SUCCESS; }
int MPI_Sendrecv(const void* sb, int sn, MPI_Datatype st, int d, int stag, void* rb, int rn, MPI_Datatype rt, int src, int rtag, MPI_Comm c, MPI_Status* s) {
  MPI_Send(sb, sn, st, d, stag, c); return MPI_Recv(rb, rn, rt, src, rtag, c, s); }
int MPI_Wait(MPI_Request* r, MPI_Status* s) {
  if (*r == 2 && g_irecv_buf && g_box) { long m = g_irecv_n < g_box_n ? g_irecv_n : g_box_n; memcpy(g_irecv_buf, g_box, m); if (s) { s->MPI_SOURCE = 0; s->MPI_TAG = g_box_tag; s->MPI_ERROR = 0; s->count_ = m; } g_irecv_buf = 0; }
  *r = MPI_REQUEST_NULL; return MPI_SUCCESS; }
int MPI_Waitall(int n, MPI_Request* r, MPI_Status* s) { for (int i = 0; i < n; ++i) MPI_Wait(r + i, s ? s + i : 0); return MPI_SUCCESS; }
int MPI_Waitsome(int n, MPI_Request* r, int* out, int* idx, MPI_Status* s) { int k = 0; for (int i = 0; i < n; ++i) if (r[i] != MPI_REQUEST_NULL) { MPI_Wait(r + i, s ? s + k : 0); idx[k++] = i; } *out = k ? k : MPI_UNDEFINED; return MPI_SUCCESS; }
int MPI_Testall(int n, MPI_Request* r, int* flag, MPI_Status* s) { MPI_Waitall(n, r, s); *flag = 1; return MPI_SUCCESS; }
int MPI_Testsome(int n, MPI_Request* r, int* out, int* idx, MPI_Status* s) { return MPI_Waitsome(n, r, out, idx, s); }
int MPI_Request_free(MPI_Request* r) { *r = MPI_REQUEST_NULL; return MPI_SUCCESS; }
int MPI_Probe(int src, int tag, MPI_Comm c, MPI_Status* s) { (void)src; (void)tag; (void)c; if (s) { s->MPI_SOURCE = 0; s->MPI_TAG = g_box_tag; s->MPI_ERROR = 0; s->count_ = g_box_n; } return MPI_SUCCESS; }
int MPI_Mprobe(int src, int tag, MPI_Comm c, MPI_Message* m, MPI_Status* s) { *m = 1; return MPI_Probe(src, tag, c, s); }
int MPI_Mrecv(void* b, int n, MPI_Datatype t, MPI_Message* m, MPI_Status* s) { *m = 0; return MPI_Recv(b, n, t, 0, 0, MPI_COMM_WORLD, s); }

struct kmi_mpi_file { int fd; };
int MPI_File_open(MPI_Comm c, const char* name, int mode, MPI_Info i, MPI_File* f) {
  (void)c; (void)mode; (void)i; int fd = open(name, O_RDONLY); if (fd < 0) { *f = MPI_FILE_NULL; return MPI_ERR_OTHER; }
  *f = (MPI_File)malloc(sizeof(struct kmi_mpi_file)); (*f)->fd = fd; return MPI_SUCCESS; }
int MPI_File_close(MPI_File* f) { if (*f) { close((*f)->fd); free(*f); } *f = MPI_FILE_NULL; return MPI_SUCCESS; }
int MPI_File_get_size(MPI_File f, MPI_Offset* s) { struct stat st; if (fstat(f->fd, &st)) return MPI_ERR_OTHER; *s = st.st_size; return MPI_SUCCESS; }
int MPI_File_set_atomicity(MPI_File f, int a) { (void)f; (void)a; return MPI_SUCCESS; }
int MPI_File_read_at(MPI_File f, MPI_Offset off, void* b, int n, MPI_Datatype t, MPI_Status* s) {
  long want = n * t->size, got = 0; while (got < want) { ssize_t r = pread(f->fd, (char*)b + got, want - got, off + got); if (r <= 0) break; got += r; }
  if (s) { s->MPI_SOURCE = 0; s->MPI_TAG = 0; s->MPI_ERROR = 0; s->count_ = got; } return MPI_SUCCESS; }
int MPI_File_read_at_all(MPI_File f, MPI_Offset off, void* b, int n, MPI_Datatype t, MPI_Status* s) { return MPI_File_read_at(f, off, b, n, t, s); }
