// fbxsdk.h shim — TEST INFRASTRUCTURE. The Autodesk FBX SDK is proprietary and
// absent; shapes/MeshObject.h:3 includes it and only names these two types.
#pragma once
class FbxMatrix;
class FbxVector4;
