// optix_stubs.h — empty on purpose (host function table of the real SDK).
#pragma once
