// TEST INFRASTRUCTURE - shared by oracle/ref_harness.cpp and integration/shim_test.cpp.
//
// ---------------------------------------------------------------------------------------------------------
// Null renderer: the reference constructors want an active renderer (ParticleEmitter fetches shaders and a
// drawable, Source/Scene/RNParticleEmitter.cpp:38-54; Mesh wants buffers).  No graphics API is involved.
// ---------------------------------------------------------------------------------------------------------
#pragma once
#include "Rayne.h"
#include <cstdlib>

namespace
{
	class NullBuffer : public GPUBuffer
	{
	public:
		NullBuffer(size_t length) : _length(length) { _data = calloc(1, length ? length : 1); }
		~NullBuffer() { free(_data); }
		void *GetBuffer() override { return _data; }
		void UnmapBuffer() override {}
		void InvalidateRange(const Range &) override {}
		void FlushRange(const Range &) override {}
		size_t GetLength() const override { return _length; }
	private:
		void *_data;
		size_t _length;
	};

	class NullShader : public Shader
	{
	public:
		NullShader(ShaderLibrary *lib, Type type, const Shader::Options *options) : Shader(lib, type, false, options) {}
		const String *GetName() const override { return RNCSTR("null"); }
	};

	class NullShaderLibrary : public ShaderLibrary
	{
	public:
		Shader *GetShaderWithName(const String *, const Shader::Options *options) override
		{
			const Shader::Options *opts = options ? options : Shader::Options::WithNone();
			return (new NullShader(this, Shader::Type::Vertex, opts))->Autorelease();
		}
		Shader *GetInstancedShaderForShader(Shader *shader) override { return shader; }
	};

	class NullRenderer : public Renderer
	{
	public:
		NullRenderer() : Renderer(reinterpret_cast<RendererDescriptor *>(16), reinterpret_cast<RenderingDevice *>(16))
		{
			_library = new NullShaderLibrary();
		}
		Window *CreateAWindow(const Vector2 &, Screen *, const Window::SwapChainDescriptor &) override { return nullptr; }
		Window *GetMainWindow() override { return nullptr; }
		void SetMainWindow(Window *) override {}
		void Render(Function &&function) override { function(); }
		void SubmitCamera(Camera *camera, Function &&function) override { current = camera; function(); current = nullptr; }
		bool SupportsTextureFormat(const String *) const override { return false; }
		bool SupportsDrawMode(DrawMode) const override { return true; }
		size_t GetAlignmentForType(PrimitiveType type) const override { return GetSizeForType(type); }
		size_t GetSizeForType(PrimitiveType type) const override
		{
			switch(type)
			{
				case PrimitiveType::Uint8: case PrimitiveType::Int8: return 1;
				case PrimitiveType::Uint16: case PrimitiveType::Int16: case PrimitiveType::Half: return 2;
				case PrimitiveType::Uint32: case PrimitiveType::Int32: case PrimitiveType::Float: return 4;
				case PrimitiveType::Vector2: return 8;
				case PrimitiveType::Vector3: return 12;
				case PrimitiveType::Vector4: case PrimitiveType::Color: case PrimitiveType::Quaternion: return 16;
				case PrimitiveType::Matrix4x4: return 64;
				default: return 16;
			}
		}
		GPUBuffer *CreateBufferWithLength(size_t length, GPUResource::UsageOptions, GPUResource::AccessOptions, bool) override
		{
			return new NullBuffer(length);
		}
		ShaderLibrary *CreateShaderLibraryWithFile(const String *) override { return _library; }
		ShaderLibrary *CreateShaderLibraryWithSource(const String *) override { return _library; }
		Shader *GetDefaultShader(Shader::Type type, Shader::Options *options, Shader::UsageHint) override
		{
			const Shader::Options *opts = options ? options : Shader::Options::WithNone();
			return (new NullShader(_library, type, opts))->Autorelease();
		}
		ShaderLibrary *GetDefaultShaderLibrary() override { return _library; }
		Texture *CreateTextureWithDescriptor(const Texture::Descriptor &) override { return nullptr; }
		Framebuffer *CreateFramebuffer(const Vector2 &) override { return nullptr; }
		Drawable *CreateDrawable() override { return new Drawable(); }
		void DeleteDrawable(Drawable *drawable) override { delete drawable; }
		void SubmitDrawable(Drawable *) override {}
		void SubmitLight(const Light *) override {}

		Camera *current = nullptr;
	private:
		NullShaderLibrary *_library;
	};
}
