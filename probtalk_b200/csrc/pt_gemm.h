// internal alias: the descriptor and activation codes live in the public C header
#pragma once
#include "../../include/probtalk_b200.h"
