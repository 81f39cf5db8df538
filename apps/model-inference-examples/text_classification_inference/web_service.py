"""The web service of the reference's example (WebServiceController.java / WebServiceDriver.java, a Spring Boot application):
`GET /` -> "welcome!", `GET /greetings?name=` -> "hello, <name>", `POST /predict` with the text as the body -> predicted class and
probability.  Here a FastAPI application; `python web_service.py` serves it with uvicorn on $PORT (8080).

    EMBEDDING_FILE_PATH=./glove.6B.300d.txt MODEL_PATH=./models/text-classification.pt python web_service.py
    curl -d "hello world, it's a nice day" localhost:8080/predict
"""
import os

from fastapi import FastAPI, Request
from fastapi.responses import PlainTextResponse
from text_classification_model import TextClassificationModel


def create_app(embedding_file_path=None, model_path=None, sequence_length=None, stop_words_count=10):
    embedding_file_path = embedding_file_path or os.environ.get("EMBEDDING_FILE_PATH", "./glove.6B.300d.txt")
    model_path = model_path or os.environ.get("MODEL_PATH", "./models/text-classification.pt")
    model = TextClassificationModel(10, stop_words_count, sequence_length or int(os.environ.get("SEQUENCE_LENGTH", 200)), embedding_file_path)
    model.loadPyTorch(model_path) if model_path.endswith((".pt", ".pth")) else model.load(model_path)
    app = FastAPI(title="text-classification-inference")

    @app.get("/", response_class=PlainTextResponse)
    def welcome():
        return "welcome!"

    @app.get("/greetings", response_class=PlainTextResponse)
    def greetings(name: str = "World"):
        return "hello, " + name

    @app.post("/predict", response_class=PlainTextResponse)
    async def predict(request: Request):
        text = (await request.body()).decode("utf-8", "replace")
        if not text:
            return "error,no text found"
        results = model.predict([[model.preprocess(text)]])
        data = results[0][0].to_ndarray().reshape(-1)
        classed = int(data.argmax())
        return "The predict class is:%s\nThe probability is:%s \n" % (classed, float(data[classed]))

    return app


if __name__ == "__main__":
    import uvicorn

    uvicorn.run(create_app(), host=os.environ.get("HOST", "127.0.0.1"), port=int(os.environ.get("PORT", 8080)))
