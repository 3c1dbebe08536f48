// Forwarder: the whole cpprl:: drop-in surface is declared in cpprl/b200/host_api.h.
#pragma once
#include "cpprl/b200/host_api.h"
