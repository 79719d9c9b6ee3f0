#pragma once
#include "ImfRgba.h"
