// Linux shim for MSVC <crtdbg.h> (RayTracingWeekend.cpp:14, :197). Test infrastructure only.
#pragma once
#define _CRTDBG_ALLOC_MEM_DF 0
#define _CRTDBG_LEAK_CHECK_DF 0
#define _CrtSetDbgFlag(x) ((void)0)
