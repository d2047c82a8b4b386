#pragma once
#include "ck/library/tensor_operation_instance/device_operation_instance_factory.hpp"
