/* TEST INFRASTRUCTURE — a single-rank stand-in for <mpi.h>, just enough for the REFERENCE's own headers (ext/mxx, src/io,
 * src/containers, src/index under /root/reference) to compile and run at p = 1, where every collective is a copy. It lets
 * oracle/ref_t1.cpp run the reference's real FASTQParser / FASTAParser / KmerParser / Index code (oracle tier
 * T1) to produce golden vectors; nothing in the product links or includes it. Not an MPI implementation: one process,
 * communicator size 1, point-to-point to self unsupported (never reached at p = 1). */
#ifndef KMI_MPI_SHIM_H_
#define KMI_MPI_SHIM_H_
#include <stddef.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MPI_VERSION 3
#define MPI_SUBVERSION 1
typedef long MPI_Aint;
typedef long long MPI_Offset;
typedef long long MPI_Count;
typedef int MPI_Comm;
typedef int MPI_Op;
typedef int MPI_Request;
typedef int MPI_Errhandler;
typedef int MPI_Info;
typedef int MPI_Message;
typedef struct kmi_mpi_file* MPI_File;
typedef struct { int MPI_SOURCE, MPI_TAG, MPI_ERROR; long long count_; } MPI_Status;
/* a datatype is its size and extent in bytes (+ lower bound, attributes) */
typedef struct kmi_mpi_type { long size, extent, lb; int builtin; void* attr[8]; int attr_set[8]; }* MPI_Datatype;
typedef void(MPI_User_function)(void*, void*, int*, MPI_Datatype*);
typedef void(MPI_Comm_errhandler_function)(MPI_Comm*, int*, ...);
typedef MPI_Comm_errhandler_function MPI_Comm_errhandler_fn;
typedef void(MPI_Handler_function)(MPI_Comm*, int*, ...);
typedef int(MPI_Type_copy_attr_function)(MPI_Datatype, int, void*, void*, void*, int*);
typedef int(MPI_Type_delete_attr_function)(MPI_Datatype, int, void*, void*);

#define MPI_SUCCESS 0
#define MPI_ERR_OTHER 15
#define MPI_COMM_NULL 0
#define MPI_COMM_WORLD 1
#define MPI_COMM_SELF 2
#define MPI_REQUEST_NULL 0
#define MPI_OP_NULL 0
#define MPI_INFO_NULL 0
#define MPI_FILE_NULL ((MPI_File)0)
#define MPI_DATATYPE_NULL ((MPI_Datatype)0)
#define MPI_STATUS_IGNORE ((MPI_Status*)0)
#define MPI_STATUSES_IGNORE ((MPI_Status*)0)
#define MPI_IN_PLACE ((void*)1)
#define MPI_UNDEFINED (-32766)
#define MPI_ANY_SOURCE (-2)
#define MPI_ANY_TAG (-1)
#define MPI_MAX_PROCESSOR_NAME 256
#define MPI_MAX_ERROR_STRING 512
#define MPI_COMM_TYPE_SHARED 1
#define MPI_MODE_RDONLY 2
#define MPI_ERRORS_RETURN 1
#define MPI_ERRORS_ARE_FATAL 2
#define MPI_TYPE_NULL_COPY_FN ((MPI_Type_copy_attr_function*)0)
#define MPI_TYPE_NULL_DELETE_FN ((MPI_Type_delete_attr_function*)0)
#define MPI_TYPE_DUP_FN ((MPI_Type_copy_attr_function*)0)
enum { MPI_SUM = 1, MPI_PROD, MPI_MAX, MPI_MIN, MPI_LAND, MPI_LOR, MPI_BAND, MPI_BOR, MPI_BXOR, MPI_LXOR, MPI_MAXLOC, MPI_MINLOC, MPI_OP_USER_ = 100 };

extern struct kmi_mpi_type kmi_mpi_builtin[32];
#define KMI_MPI_T(i) (&kmi_mpi_builtin[i])
#define MPI_CHAR KMI_MPI_T(0)
#define MPI_SIGNED_CHAR KMI_MPI_T(1)
#define MPI_UNSIGNED_CHAR KMI_MPI_T(2)
#define MPI_BYTE KMI_MPI_T(3)
#define MPI_SHORT KMI_MPI_T(4)
#define MPI_UNSIGNED_SHORT KMI_MPI_T(5)
#define MPI_INT KMI_MPI_T(6)
#define MPI_UNSIGNED KMI_MPI_T(7)
#define MPI_LONG KMI_MPI_T(8)
#define MPI_UNSIGNED_LONG KMI_MPI_T(9)
#define MPI_LONG_LONG KMI_MPI_T(10)
#define MPI_LONG_LONG_INT KMI_MPI_T(10)
#define MPI_UNSIGNED_LONG_LONG KMI_MPI_T(11)
#define MPI_FLOAT KMI_MPI_T(12)
#define MPI_DOUBLE KMI_MPI_T(13)
#define MPI_LONG_DOUBLE KMI_MPI_T(14)
#define MPI_FLOAT_INT KMI_MPI_T(15)
#define MPI_DOUBLE_INT KMI_MPI_T(16)
#define MPI_LONG_INT KMI_MPI_T(17)
#define MPI_2INT KMI_MPI_T(18)
#define MPI_SHORT_INT KMI_MPI_T(19)
#define MPI_LONG_DOUBLE_INT KMI_MPI_T(20)
#define MPI_WCHAR KMI_MPI_T(21)
#define MPI_C_BOOL KMI_MPI_T(22)
#define MPI_INT8_T KMI_MPI_T(1)
#define MPI_UINT8_T KMI_MPI_T(2)
#define MPI_INT16_T KMI_MPI_T(4)
#define MPI_UINT16_T KMI_MPI_T(5)
#define MPI_INT32_T KMI_MPI_T(6)
#define MPI_UINT32_T KMI_MPI_T(7)
#define MPI_INT64_T KMI_MPI_T(8)
#define MPI_UINT64_T KMI_MPI_T(9)

int MPI_Init(int*, char***);
int MPI_Init_thread(int*, char***, int, int*);
int MPI_Finalize(void);
int MPI_Initialized(int*);
int MPI_Finalized(int*);
int MPI_Abort(MPI_Comm, int);
int MPI_Comm_rank(MPI_Comm, int*);
int MPI_Comm_size(MPI_Comm, int*);
int MPI_Comm_dup(MPI_Comm, MPI_Comm*);
int MPI_Comm_free(MPI_Comm*);
int MPI_Comm_split(MPI_Comm, int, int, MPI_Comm*);
int MPI_Comm_split_type(MPI_Comm, int, int, MPI_Info, MPI_Comm*);
int MPI_Comm_compare(MPI_Comm, MPI_Comm, int*);
int MPI_Comm_set_errhandler(MPI_Comm, MPI_Errhandler);
int MPI_Comm_create_errhandler(MPI_Comm_errhandler_function*, MPI_Errhandler*);
int MPI_Errhandler_create(MPI_Handler_function*, MPI_Errhandler*);
int MPI_Errhandler_set(MPI_Comm, MPI_Errhandler);
int MPI_Errhandler_free(MPI_Errhandler*);
int MPI_Error_string(int, char*, int*);
int MPI_Error_class(int, int*);
int MPI_Get_processor_name(char*, int*);
int MPI_Barrier(MPI_Comm);
double MPI_Wtime(void);
/* datatypes */
int MPI_Type_contiguous(int, MPI_Datatype, MPI_Datatype*);
int MPI_Type_vector(int, int, int, MPI_Datatype, MPI_Datatype*);
int MPI_Type_create_struct(int, const int*, const MPI_Aint*, const MPI_Datatype*, MPI_Datatype*);
int MPI_Type_create_resized(MPI_Datatype, MPI_Aint, MPI_Aint, MPI_Datatype*);
int MPI_Type_dup(MPI_Datatype, MPI_Datatype*);
int MPI_Type_commit(MPI_Datatype*);
int MPI_Type_free(MPI_Datatype*);
int MPI_Type_size(MPI_Datatype, int*);
int MPI_Type_get_extent(MPI_Datatype, MPI_Aint*, MPI_Aint*);
int MPI_Type_get_true_extent(MPI_Datatype, MPI_Aint*, MPI_Aint*);
int MPI_Type_create_keyval(MPI_Type_copy_attr_function*, MPI_Type_delete_attr_function*, int*, void*);
int MPI_Type_set_attr(MPI_Datatype, int, void*);
int MPI_Type_get_attr(MPI_Datatype, int, void*, int*);
int MPI_Get_address(const void*, MPI_Aint*);
int MPI_Get_count(const MPI_Status*, MPI_Datatype, int*);
int MPI_Get_elements_x(const MPI_Status*, MPI_Datatype, MPI_Count*);
int MPI_Op_create(MPI_User_function*, int, MPI_Op*);
int MPI_Op_free(MPI_Op*);
/* collectives: copies at p = 1 */
int MPI_Bcast(void*, int, MPI_Datatype, int, MPI_Comm);
int MPI_Reduce(const void*, void*, int, MPI_Datatype, MPI_Op, int, MPI_Comm);
int MPI_Allreduce(const void*, void*, int, MPI_Datatype, MPI_Op, MPI_Comm);
int MPI_Scan(const void*, void*, int, MPI_Datatype, MPI_Op, MPI_Comm);
int MPI_Exscan(const void*, void*, int, MPI_Datatype, MPI_Op, MPI_Comm);
int MPI_Gather(const void*, int, MPI_Datatype, void*, int, MPI_Datatype, int, MPI_Comm);
int MPI_Gatherv(const void*, int, MPI_Datatype, void*, const int*, const int*, MPI_Datatype, int, MPI_Comm);
int MPI_Allgather(const void*, int, MPI_Datatype, void*, int, MPI_Datatype, MPI_Comm);
int MPI_Allgatherv(const void*, int, MPI_Datatype, void*, const int*, const int*, MPI_Datatype, MPI_Comm);
int MPI_Scatter(const void*, int, MPI_Datatype, void*, int, MPI_Datatype, int, MPI_Comm);
int MPI_Scatterv(const void*, const int*, const int*, MPI_Datatype, void*, int, MPI_Datatype, int, MPI_Comm);
int MPI_Alltoall(const void*, int, MPI_Datatype, void*, int, MPI_Datatype, MPI_Comm);
int MPI_Alltoallv(const void*, const int*, const int*, MPI_Datatype, void*, const int*, const int*, MPI_Datatype, MPI_Comm);
int MPI_Alltoallw(const void*, const int*, const int*, const MPI_Datatype*, void*, const int*, const int*, const MPI_Datatype*, MPI_Comm);
/* point to point: only a matching send/recv to self through a one-message mailbox */
int MPI_Send(const void*, int, MPI_Datatype, int, int, MPI_Comm);
int MPI_Recv(void*, int, MPI_Datatype, int, int, MPI_Comm, MPI_Status*);
int MPI_Isend(const void*, int, MPI_Datatype, int, int, MPI_Comm, MPI_Request*);
int MPI_Irecv(void*, int, MPI_Datatype, int, int, MPI_Comm, MPI_Request*);
int MPI_Sendrecv(const void*, int, MPI_Datatype, int, int, void*, int, MPI_Datatype, int, int, MPI_Comm, MPI_Status*);
int MPI_Wait(MPI_Request*, MPI_Status*);
int MPI_Waitall(int, MPI_Request*, MPI_Status*);
int MPI_Waitsome(int, MPI_Request*, int*, int*, MPI_Status*);
int MPI_Testall(int, MPI_Request*, int*, MPI_Status*);
int MPI_Testsome(int, MPI_Request*, int*, int*, MPI_Status*);
int MPI_Request_free(MPI_Request*);
int MPI_Probe(int, int, MPI_Comm, MPI_Status*);
int MPI_Mprobe(int, int, MPI_Comm, MPI_Message*, MPI_Status*);
int MPI_Mrecv(void*, int, MPI_Datatype, MPI_Message*, MPI_Status*);
/* MPI-IO: plain pread */
int MPI_File_open(MPI_Comm, const char*, int, MPI_Info, MPI_File*);
int MPI_File_close(MPI_File*);
int MPI_File_get_size(MPI_File, MPI_Offset*);
int MPI_File_set_atomicity(MPI_File, int);
int MPI_File_read_at_all(MPI_File, MPI_Offset, void*, int, MPI_Datatype, MPI_Status*);
int MPI_File_read_at(MPI_File, MPI_Offset, void*, int, MPI_Datatype, MPI_Status*);

#ifdef __cplusplus
}
#endif
#endif
