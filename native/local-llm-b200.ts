// local-llm-b200.ts -- the adapter class a maintainer drops into src/adapters/ of the reference: the same contract as
// LocalLlmAdapter (src/adapters/local-llm.ts:6-22, base.ts:10-29), calling the B200 engine in-process through the
// Node-API addon (native/rt_addon.c) instead of fetch().  NOT COMPILED on the build image (no Node, no tsc); kept as a
// file so that it can be type-checked and built where the reference's toolchain exists (Node >= 20, package.json:53-55).
//
// Registration: one branch in createAdapter (src/utils/adapters.ts:47-51) -- when the `local-llm*` entry's endpoint
// starts with "b200://", construct LocalLlmB200Adapter(endpoint, model, name ?? id, source, timeoutMs) instead of
// LocalLlmAdapter.  theroundtaible_b200/adapter.py::create_adapter is the same rule in Python.
import { BaseAdapter } from "./base.js";
import { classifyError } from "../utils/errors.js";
import { createRequire } from "node:module";

interface NativeAddon {
  create(model: string, device: number, maxCtx: number, maxNewTokens?: number): unknown;
  generate(engine: unknown, prompt: string, timeoutMs: number): Promise<string>;
  contextLength(engine: unknown): number;
  destroy(engine: unknown): void;
}

const native = createRequire(import.meta.url)("../../native/rt_addon.node") as NativeAddon;

const MAX_CTX = 65536;          // byte-level tokenizer: one token per UTF-8 byte, so windows are sized in chars
const MAX_NEW_TOKENS = 512;     // templates/system-prompt.md:17 asks for at most 500 words

export class LocalLlmB200Adapter extends BaseAdapter {
  readonly name: string;

  private endpoint: string;
  private model: string;
  private defaultTimeout: number;
  private engine: unknown;

  constructor(endpoint: string, model: string, name: string, _source?: string, timeoutMs: number = 120_000) {
    super();
    this.endpoint = endpoint.replace(/\/+$/, "");   // "b200://<device ordinal>"
    this.model = model;
    this.name = name;
    this.defaultTimeout = timeoutMs;
  }

  private device(): number {
    const m = /^b200:\/\/(\d+)$/.exec(this.endpoint);
    if (!m) throw new Error(`endpoint '${this.endpoint}' not found: expected b200://<device ordinal>`);
    return Number(m[1]);
  }

  // local-llm.ts:24-36: true when the backend answers; never throws
  async isAvailable(): Promise<boolean> {
    try {
      this.engine ??= native.create(this.model, this.device(), MAX_CTX, MAX_NEW_TOKENS);
      return true;
    } catch {
      return false;
    }
  }

  // local-llm.ts:58-70 reserves 4096 + 3000 tokens and assumes 4 chars per token; here a char IS a token
  override getMaxSourceChars(): number | undefined {
    if (!this.engine) return undefined;
    const ctx = native.contextLength(this.engine);
    return Math.max(ctx - MAX_NEW_TOKENS - 3000 * 4, 2000);
  }

  // local-llm.ts:72-89,150-199: the generated text, or a rejection classified by classifyError
  // ("... timed out ..." -> kind "timeout"; "context length exceeded" is what isContextWindowError looks for)
  async execute(prompt: string, timeoutMs?: number): Promise<string> {
    try {
      if (!this.engine && !(await this.isAvailable())) {
        throw new Error(`model '${this.model}' not found or no B200 at ${this.endpoint}`);
      }
      const text = await native.generate(this.engine, prompt, timeoutMs ?? this.defaultTimeout);
      if (!text) throw new Error("Local LLM returned empty response");          // local-llm.ts:189-191
      return text;
    } catch (error) {
      throw classifyError(error, this.name);
    }
  }
}
