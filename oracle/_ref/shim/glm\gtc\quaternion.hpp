#include <glm/gtc/quaternion.hpp>
