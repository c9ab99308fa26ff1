/* Single-process stand-in for <mpi.h>, test infrastructure only.
 * mini/mesh/part.hpp names these symbols (part.hpp:24,623,641,911,926,1271,1280) but the
 * oracle harness never instantiates mesh::part::Part itself, so the calls are never
 * executed: they only have to parse.  Not used by the product. */
#ifndef ORACLE_STUB_MPI_H_
#define ORACLE_STUB_MPI_H_
typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Request;
typedef int MPI_Info;
typedef struct { int unused; } MPI_Status;
#define MPI_COMM_WORLD 0
#define MPI_LONG 1
#define MPI_INT 2
#define MPI_DOUBLE 3
#define MPI_FLOAT 4
#define MPI_INFO_NULL 0
static inline int MPI_Isend(const void *, int, MPI_Datatype, int, int, MPI_Comm, MPI_Request *) { return 0; }
static inline int MPI_Irecv(void *, int, MPI_Datatype, int, int, MPI_Comm, MPI_Request *) { return 0; }
static inline int MPI_Waitall(int, MPI_Request *, MPI_Status *) { return 0; }
static inline int MPI_Barrier(MPI_Comm) { return 0; }
#endif
