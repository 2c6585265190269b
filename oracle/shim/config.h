/*
 * Hand-written stand-in for FFmpeg's configure-generated config.h.
 * TEST INFRASTRUCTURE ONLY: lets oracle/Makefile compile the reference's own VVC
 * reconstruction sources (read in place from /root/reference) as plain C with no
 * assembler, no threads and no external libraries.  Nothing here is product code.
 */
#ifndef ORACLE_SHIM_CONFIG_H
#define ORACLE_SHIM_CONFIG_H
#define FFMPEG_CONFIGURATION "oracle-shim"
#define FFMPEG_LICENSE "LGPL version 2.1 or later"
#define av_restrict restrict
#define EXTERN_PREFIX ""
#define EXTERN_ASM
#define BUILDSUF ""
#define SLIBSUF ".so"
#define SWS_MAX_FILTER_SIZE 256

#define ARCH_AARCH64 0
#define ARCH_ALPHA 0
#define ARCH_ARM 0
#define ARCH_AVR32 0
#define ARCH_LOONGARCH 0
#define ARCH_LOONGARCH32 0
#define ARCH_LOONGARCH64 0
#define ARCH_MIPS 0
#define ARCH_MIPS64 0
#define ARCH_PPC 0
#define ARCH_PPC64 0
#define ARCH_RISCV 0
#define ARCH_S390 0
#define ARCH_SH4 0
#define ARCH_SPARC 0
#define ARCH_X86 0
#define ARCH_X86_32 0
#define ARCH_X86_64 0

#define HAVE_BIGENDIAN 0
#define HAVE_FAST_UNALIGNED 1
#define HAVE_FAST_64BIT 1
#define HAVE_FAST_CLZ 1
#define HAVE_FAST_CMOV 0
#define HAVE_LOCAL_ALIGNED 1
#define HAVE_SIMD_ALIGN_16 0
#define HAVE_SIMD_ALIGN_32 0
#define HAVE_SIMD_ALIGN_64 0
#define HAVE_INLINE_ASM 0
#define HAVE_INLINE_ASM_DIRECT_SYMBOL_REFS 0
#define HAVE_X86ASM 0
#define HAVE_MMX 0
#define HAVE_MMX_INLINE 0
#define HAVE_MMX_EXTERNAL 0
#define HAVE_MIPSFPU 0
#define HAVE_6REGS 0
#define HAVE_7REGS 0
#define HAVE_EBP_AVAILABLE 0
#define HAVE_EBX_AVAILABLE 0
#define HAVE_XMM_CLOBBERS 0
#define HAVE_RDTSC 0
#define HAVE_LIBC_MSVCRT 0
#define HAVE_PRAGMA_DEPRECATED 1
#define HAVE_THREADS 0
#define HAVE_PTHREADS 0
#define HAVE_W32THREADS 0
#define HAVE_OS2THREADS 0
#define HAVE_ATOMICS_NATIVE 1

#define HAVE_ATAN2F 1
#define HAVE_ATANF 1
#define HAVE_CBRT 1
#define HAVE_CBRTF 1
#define HAVE_COPYSIGN 1
#define HAVE_COSF 1
#define HAVE_ERF 1
#define HAVE_EXP2 1
#define HAVE_EXP2F 1
#define HAVE_EXPF 1
#define HAVE_HYPOT 1
#define HAVE_ISFINITE 1
#define HAVE_ISINF 1
#define HAVE_ISNAN 1
#define HAVE_LDEXPF 1
#define HAVE_LLRINT 1
#define HAVE_LLRINTF 1
#define HAVE_LOG10F 1
#define HAVE_LOG2 1
#define HAVE_LOG2F 1
#define HAVE_LRINT 1
#define HAVE_LRINTF 1
#define HAVE_POWF 1
#define HAVE_RINT 1
#define HAVE_ROUND 1
#define HAVE_ROUNDF 1
#define HAVE_SINF 1
#define HAVE_TRUNC 1
#define HAVE_TRUNCF 1

#define CONFIG_SMALL 0
#define CONFIG_SHARED 0
#define CONFIG_GRAY 0
#define CONFIG_MEMORY_POISONING 0
#define CONFIG_SAFE_BITSTREAM_READER 1
#define CONFIG_VVC_DECODER 1
#endif
