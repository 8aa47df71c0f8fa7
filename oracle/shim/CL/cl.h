/* TEST INFRASTRUCTURE ONLY. Empty stand-in for <CL/cl.h>: the reference's host data-model
 * headers include it but the BVH builder uses nothing from it. No OpenCL runtime exists on
 * this box. */
#ifndef ORACLE_STUB_CL_H
#define ORACLE_STUB_CL_H
#endif
