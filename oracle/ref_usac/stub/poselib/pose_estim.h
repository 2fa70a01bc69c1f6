#pragma once
#include "pose_helper.h"
