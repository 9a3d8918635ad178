// Stand-in for shaders/common_bindings.h when the shaders are compiled as C++ (oracle/_ref). The reference file holds
// only descriptor declarations (bindings 0-5); here the same names are plain members of the per-shader class.
// Field names and types follow shaders/common_bindings.h:1-34; included inside a class body, after RayLight.
image2D img;                         // binding 0 (rgba8 storage image)
RayLight* lights = nullptr;          // binding 3
uint lightCount = 0;                 // binding 4
uint contributionCounter = 0;
// binding 5 (UBO)
uint frameCount = 0;
uvec2 framebufferDim;
vec3 cameraPos;
float fov = 0;
float photonInitialIntensity = 0;
float scatteringProbability = 0;
float absorptionCoefficient = 0;
uint maxLights = 0;
float rayMaxDistance = 0;
float rayMarchingStepSize = 0;
vec3 lightSourceWorldPos;
float beamRadius = 0;
float lightRayStepSize = 0;
float radiusFalloff = 0;
// compute built-in
uvec3 gl_GlobalInvocationID;
