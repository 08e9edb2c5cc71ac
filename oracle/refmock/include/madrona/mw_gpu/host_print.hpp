/* shim: mwGPU::HostPrint::log (device printf ring, M/src/mw/device/host_print.cpp) -> nothing */
#pragma once
namespace madrona { namespace mwGPU { struct HostPrint { template <typename... A> static void log(const char *, A...) {} }; } }
