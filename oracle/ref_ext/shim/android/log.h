// stands in for <android/log.h>: the reference's timing log lines compile to nothing (oracle/ref_ext, test infrastructure)
#ifndef ORBFE_ORACLE_ANDROID_LOG_SHIM_H
#define ORBFE_ORACLE_ANDROID_LOG_SHIM_H
#define ANDROID_LOG_INFO 4
#define __android_log_print(...) ((void)0)
#endif
